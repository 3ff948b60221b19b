"""Where the end-to-end time goes: gjd_decode_batch (host buffers) under different sub-batch sizes and host
thread counts, next to the raw pinned D2H bandwidth of the same number of bytes.
usage: python tools/e2e_probe.py [images]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth
import gjd
import torch

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
base = gen_synth.c3_files()
files = [base[i % len(base)] for i in range(n)]
out_bytes = 3 * 1024 * 1024
pinned = gjd.PinnedBuffer(out_bytes * n)

# raw D2H of the same bytes from a device buffer into the same pinned memory
dev = torch.empty(out_bytes * n, dtype=torch.uint8, device="cuda")
host = torch.frombuffer(pinned.array, dtype=torch.uint8)
for _ in range(2):
    t0 = time.perf_counter(); host.copy_(dev, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("raw D2H %.1f GB in %.1f ms = %.1f GB/s" % (out_bytes * n / 1e9, dt * 1e3, out_bytes * n / dt / 1e9), flush=True)
del dev

for scratch_gb, threads in ((4, 0), (2, 0), (1, 0), (8, 0), (4, 8), (4, 32)):
    os.environ["GJD_SCRATCH"] = str(scratch_gb << 30)
    if threads:
        os.environ["GJD_HOST_THREADS"] = str(threads)
    else:
        os.environ.pop("GJD_HOST_THREADS", None)
    with gjd.Decoder(0, scratch_bytes=scratch_gb << 30) as d:
        items, hold = d.make_items(files)
        for i in range(n):
            items[i].rgb = pinned.ptr + i * out_bytes
        d.decode_batch_raw(items, n)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); d.decode_batch_raw(items, n); ts.append(time.perf_counter() - t0)
        print("scratch %d GB threads %s: e2e %.1f ms (min of 3) = %.1f GPix/s" %
              (scratch_gb, threads or "default", min(ts) * 1e3, n * 1.048576 / min(ts) / 1e3), flush=True)
