"""The other BASELINE.json configs on one GPU (bench.py times config 3 only):
  C1  cuda-decoder/profile.jpg (= tests/golden/images/2_400x400.jpg): single-image latency through gjd_decode_batch
  C4  one 8192x8192 q90 image without restart markers: stage times of the intra-image parallel decode
  C5  mixed sizes (256..2048 squared + 10 % ragged, q 75/90/95), 256 distinct files tiled to N images: resident
      throughput, and gjd_partition's balance over 2/4/8 parts
usage: python tools/config_bench.py [c5_images=6000] [c5_distinct=256]      prints one JSON line per config"""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth
import gjd
import numpy as np

n5 = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
d5 = int(sys.argv[2]) if len(sys.argv) > 2 else 256


def resident(dec, files, runs=10):
    b = dec.prepare(files)
    for _ in range(3):
        b.run()
        assert b.sync() == 0
    t = []
    for _ in range(runs):
        b.run()
        assert b.sync() == 0
        t.append(dec.stage_times())
    info = b.info()
    b.free()
    med = [statistics.median(x[i] for x in t) for i in range(4)]
    return med, info


with gjd.Decoder(0) as dec:
    # ---- C1 ----
    with open(os.path.join(ROOT, "tests", "golden", "images", "2_400x400.jpg"), "rb") as f:
        one = f.read()
    ts = []
    for it in range(0 if os.environ.get("CONFIG_BENCH_SKIP_C1") else 33):
        t0 = time.perf_counter()
        outs, st = dec.decode([one])
        if it >= 3:
            ts.append(time.perf_counter() - t0)
    if ts:
        med, _ = resident(dec, [one], 20)
        print(json.dumps({"config": "c1 profile.jpg 400x400", "host_call_ms_median": round(statistics.median(ts) * 1e3, 3),
                          "host_call_ms_min": round(min(ts) * 1e3, 3), "device_ms": round(med[3], 4),
                          "stages_ms": [round(x, 4) for x in med[:3]]}), flush=True)

    # ---- C4 ----
    big = gen_synth.synth_jpeg(1000, 8192, 8192, 90)
    med, info = resident(dec, [big], 10)
    print(json.dumps({"config": "c4 8192x8192 q90, no restart markers", "scan_bytes": info["scan_bytes"],
                      "subsequences": info["subseqs"], "device_ms": round(med[3], 3),
                      "stages_ms": [round(x, 3) for x in med[:3]],
                      "MPixel_per_s": round(info["pixels"] / med[3] / 1e3, 1)}), flush=True)

    # ---- C4 shape, 4:2:0: intra-image parallelism of the chroma-subsampled path (DC chains per tile, not per image) ----
    big420 = gen_synth.synth_jpeg(1000, 8192, 8192, 90, False, 0, "2x2")
    med, info = resident(dec, [big420], 10)
    print(json.dumps({"config": "c4 shape 8192x8192 q90 4:2:0, no restart markers", "scan_bytes": info["scan_bytes"],
                      "subsequences": info["subseqs"], "device_ms": round(med[3], 3),
                      "stages_ms": [round(x, 3) for x in med[:3]],
                      "MPixel_per_s": round(info["pixels"] / med[3] / 1e3, 1)}), flush=True)

    # ---- C3 with restart intervals (one MCU row per interval, the usual camera setting): ADVICE r1 ----
    rbase = [gen_synth.synth_jpeg(s, 1024, 1024, 90, False, 1) for s in range(16)]
    nbase = [gen_synth.synth_jpeg(s, 1024, 1024, 90) for s in range(16)]
    for label, base_files in (("c3 shape, 1000 x 1024x1024 q90, no restart markers", nbase),
                              ("c3 shape, 1000 x 1024x1024 q90, DRI = one MCU row (128 intervals per image)", rbase)):
        med, info = resident(dec, [base_files[i % 16] for i in range(1000)], 10)
        print(json.dumps({"config": label, "subsequences": info["subseqs"], "device_ms": round(med[3], 3),
                          "stages_ms": [round(x, 3) for x in med[:3]],
                          "MPixel_per_s": round(info["pixels"] / med[3] / 1e3, 1)}), flush=True)

    # ---- C3 shape, chroma-subsampled: the same 16 pictures encoded 4:2:2 and 4:2:0 ----
    for label, sample in (("4:2:2 (-sample 2x1)", "2x1"), ("4:2:0 (-sample 2x2)", "2x2")):
        sbase = [gen_synth.synth_jpeg(s, 1024, 1024, 90, False, 0, sample) for s in range(16)]
        med, info = resident(dec, [sbase[i % 16] for i in range(1000)], 10)
        print(json.dumps({"config": "c3 shape, 1000 x 1024x1024 q90, " + label, "subsequences": info["subseqs"],
                          "scan_bytes": info["scan_bytes"], "device_ms": round(med[3], 3), "stages_ms": [round(x, 3) for x in med[:3]],
                          "MPixel_per_s": round(info["pixels"] / med[3] / 1e3, 1)}), flush=True)

    # ---- C5 ----
    if n5 <= 0:
        sys.exit(0)
    specs = gen_synth.c5_specs(d5)
    base = [gen_synth.synth_jpeg(*s) for s in specs]
    files = [base[i % len(base)] for i in range(n5)]
    med, info = resident(dec, files, 5)
    line = {"config": "c5 mixed sizes, %d images (%d distinct)" % (n5, d5), "pixels": info["pixels"],
            "scan_bytes": info["scan_bytes"], "device_ms": round(med[3], 3), "stages_ms": [round(x, 3) for x in med[:3]],
            "MPixel_per_s": round(info["pixels"] / med[3] / 1e3, 1)}
    for parts in (2, 4, 8):
        part, load, weight = gjd.partition(files, parts)
        load = np.asarray(load, dtype=np.float64)
        line["lpt_imbalance_%d" % parts] = round(float(load.max() / load.mean() - 1.0), 5)
    print(json.dumps(line), flush=True)
