"""Small driver for ncu: prepare a resident c3 batch and run the GPU pipeline a few times.
usage: python tools/profile_run.py [images] [runs] [c3|c5|422|420|dri]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth
import gjd

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
which = sys.argv[3] if len(sys.argv) > 3 else "c3"
if which in ("422", "420"):
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90, False, 0, "2x1" if which == "422" else "2x2") for s in range(16)]
elif which == "c5":
    base = gen_synth.c5_files(256)
elif which == "dri":
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90, False, 1) for s in range(16)]
else:
    base = gen_synth.c3_files()
files = [base[i % len(base)] for i in range(n)]
with gjd.Decoder(0) as d:
    b = d.prepare(files)
    for _ in range(runs):
        b.run()
        assert b.sync() == 0
    print("stage ms", d.stage_times(), b.info())
    b.free()
