"""gjd_decode_batch with host JPEG bytes in and pixels left in DEVICE memory (rgb_on_device = 1): the
data-loader use of the library — parse + unstuff + H2D + GPU stages, no read-back.
usage: python tools/e2e_device_out.py [images] [layout 0|1]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth
import gjd
import numpy as np
import torch

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
layout = int(sys.argv[2]) if len(sys.argv) > 2 else 0
base = gen_synth.c3_files()
files = [base[i % len(base)] for i in range(n)]
out_bytes = 3 * 1024 * 1024
dev = torch.zeros(out_bytes * n, dtype=torch.uint8, device="cuda")
for scratch_gb in (1, 2, 4):
    with gjd.Decoder(0, scratch_bytes=scratch_gb << 30) as d:
        items, hold = d.make_items(files)
        for i in range(n):
            items[i].rgb = dev.data_ptr() + i * out_bytes
            items[i].rgb_on_device = 1
            items[i].layout = layout
        d.decode_batch_raw(items, n)
        ts = []
        for _ in range(5):
            t0 = time.perf_counter(); d.decode_batch_raw(items, n); ts.append(time.perf_counter() - t0)
        print("scratch %d GB: device-output e2e %.1f ms (min of 5) = %.1f GPix/s" %
              (scratch_gb, min(ts) * 1e3, n * 1.048576 / min(ts) / 1e3), flush=True)
# spot check against the host-output path
with gjd.Decoder(0) as d:
    outs, st = d.decode(files[:2], layout=layout)
got = dev[: out_bytes].cpu().numpy().reshape(outs[0].shape)
print("first image equals host-output decode:", bool(np.array_equal(got, outs[0])))
