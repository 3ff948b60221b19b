"""N ranks at once (torchrun): raw pinned D2H bandwidth per rank, then gjd_decode_batch end to end with
different host-thread counts.  Shows what bounds the end-to-end figure when the GPUs share one host.
usage: python -m torch.distributed.run --nproc-per-node N tools/e2e_probe_multi.py [images]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import torch.distributed as dist
import gen_synth
import gjd

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
base = gen_synth.c3_files()
files = [base[i % len(base)] for i in range(n)]
out_bytes = 3 * 1024 * 1024
pinned = gjd.PinnedBuffer(out_bytes * n)


def allmax(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


dev = torch.empty(out_bytes * n, dtype=torch.uint8, device="cuda")
host = torch.frombuffer(pinned.array, dtype=torch.uint8)
for _ in range(2):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter(); host.copy_(dev, non_blocking=True); torch.cuda.synchronize(); dt = allmax(time.perf_counter() - t0)
if rank == 0:
    print("raw D2H, %d ranks at once: %.1f ms = %.1f GB/s per rank, %.1f GB/s total" %
          (world, dt * 1e3, out_bytes * n / dt / 1e9, world * out_bytes * n / dt / 1e9), flush=True)
del dev
for threads in (0, 4, 2, 8):
    if threads:
        os.environ["GJD_HOST_THREADS"] = str(threads)
    with gjd.Decoder(local, scratch_bytes=4 << 30) as d:
        items, hold = d.make_items(files)
        for i in range(n):
            items[i].rgb = pinned.ptr + i * out_bytes
        d.decode_batch_raw(items, n)
        best = 1e9
        for _ in range(2):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter(); d.decode_batch_raw(items, n); dt = allmax(time.perf_counter() - t0)
            best = min(best, dt)
        if rank == 0:
            print("threads %s: e2e %.1f ms = %.1f GPix/s total" % (threads or "default", best * 1e3, world * n * 1.048576 / best / 1e3), flush=True)
dist.destroy_process_group()
