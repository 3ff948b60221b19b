"""Damaged files through libgjd, for `compute-sanitizer --tool memcheck python tools/fuzz_gpu.py [cases] [seed]`:
a damaged file must end in a status code without any out-of-bounds access on the device.
Prints the histogram of statuses.  FUZZ_BIG=1: seeds are the 1024x1024 synthetic files (4:4:4, 4:2:0, 4:2:2, with and without
restart intervals) instead of the small fixtures."""
import collections
import glob
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import fuzzlib
import gjd

n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
seeds = []
if os.environ.get("FUZZ_BIG"):
    # long scans (1024x1024): many sequences per image, checkpointed repeats, restart tables that span thread blocks
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_synth
    for sample, restart in (("1x1", 0), ("2x2", 0), ("2x1", 0), ("1x1", 1), ("2x2", 1), ("2x1", "11B")):
        seeds.append(gen_synth.synth_jpeg(7 if restart and sample != "1x1" else 0, 1024, 1024, 90, False, restart, sample))
else:
    for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*", "*.jpg"))):
        if os.path.getsize(f) < 60000:
            seeds.append(open(f, "rb").read())
files = []
for f in fuzzlib.cases(seeds, n, seed):
    rc, w, h = gjd.peek(f)
    if rc != 0 or w * h <= 1 << 20:
        files.append(f)
hist = collections.Counter()
with gjd.Decoder(0) as d:
    for first in range(0, len(files), 100):
        outs, st = d.decode(files[first : first + 100])
        hist.update(st)
print("statuses:", dict(hist), flush=True)
