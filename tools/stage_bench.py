"""A/B of kernel variants on a resident c3 batch: median CUDA-event stage times per GJD_TRANSFORM_VARIANT.
usage: python tools/stage_bench.py [images] [runs] [variants, comma separated] [c3|c5|422|420|dri = which synthetic set is tiled]"""
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth
import gjd

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 10
variants = [int(v) for v in (sys.argv[3] if len(sys.argv) > 3 else "0").split(",")]
which = sys.argv[4] if len(sys.argv) > 4 else "c3"
if which == "c5":
    base = gen_synth.c5_files(256)
elif which in ("422", "420"):    # the 16 pictures of tools/config_bench.py's subsampled lines
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90, False, 0, "2x1" if which == "422" else "2x2") for s in range(16)]
elif which == "dri":               # the restart-interval line of tools/config_bench.py (one MCU row per interval)
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90, False, 1) for s in range(16)]
else:
    base = gen_synth.c3_files()
files = [base[i % len(base)] for i in range(n)]
for v in variants:
    os.environ["GJD_TRANSFORM_VARIANT"] = str(v)
    with gjd.Decoder(0) as d:
        b = d.prepare(files)
        for _ in range(3):
            b.run()
            assert b.sync() == 0
        t = []
        for _ in range(runs):
            b.run()
            assert b.sync() == 0
            t.append(d.stage_times())
        med = [round(statistics.median(x[i] for x in t), 3) for i in range(4)]
        print("variant", v, "sync/write/transform/pipeline ms (median of %d):" % runs, med, flush=True)
        b.free()
