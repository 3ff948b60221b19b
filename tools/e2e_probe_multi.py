"""N ranks at once (torchrun; also runs as one plain process): what bounds gjd_decode_batch when the GPUs share one host.
  1. bare pinned D2H of the step's pixels and bare pinned H2D of the step's JPEG bytes, all ranks at once (platform ceilings)
  2. gjd_decode_batch with the pixels left on the device (no read-back), for each input path:
       pageable inputs + host unstuffing   (GJD_DEVICE_UNSTUFF=0: round 1's path, the host byte loop on the critical path)
       pageable inputs + device unstuffing (a memcpy into pinned staging is all the host does per byte)
       pinned inputs   + device unstuffing (DMA straight from the caller's buffer: no host pass over the bytes at all)
     and the same with the pixels read back into pinned host memory.
  Per-rank milliseconds (max over ranks) and the whole-job GPixel/s; the kernel time (gjd_stage_times) beside them.
usage: python -m torch.distributed.run --nproc-per-node N tools/e2e_probe_multi.py [images]   |   python tools/e2e_probe_multi.py [images]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gpu-jpeg-decoder_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import gen_synth
import gjd

rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
base = gen_synth.c3_files()
files = [base[i % len(base)] for i in range(n)]
out_bytes = 3 * 1024 * 1024
in_bytes = sum(len(f) for f in files)
pinned = gjd.PinnedBuffer(out_bytes * n)
if "GJD_HOST_THREADS" not in os.environ and world > 1:
    os.environ["GJD_HOST_THREADS"] = str(max(2, (os.cpu_count() or 2) // world))


def barrier():
    if dist:
        dist.barrier()
    torch.cuda.synchronize()


def allmax(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def say(*a):
    if rank == 0:
        print(*a, flush=True)


say("ranks %d, host cores %d, host threads per rank %s, %d images per rank (%.2f GB of JPEG, %.2f GB of pixels)" %
    (world, os.cpu_count(), os.environ.get("GJD_HOST_THREADS", "all"), n, in_bytes / 1e9, out_bytes * n / 1e9))

# ---- 1. bare copies -------------------------------------------------------------------------------------------------
dev = torch.empty(out_bytes * n, dtype=torch.uint8, device="cuda")
host = torch.frombuffer(pinned.array, dtype=torch.uint8)
for name, src, dst, nbytes in (("D2H of the pixels", dev, host, out_bytes * n), ("H2D of the JPEG bytes", host[:in_bytes], dev[:in_bytes], in_bytes)):
    best = 1e9
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, allmax(time.perf_counter() - t0))
    say("bare pinned %s, %d ranks at once: %.2f ms = %.1f GB/s per rank, %.1f GB/s total" %
        (name, world, best * 1e3, nbytes / best / 1e9, world * nbytes / best / 1e9))

# ---- 2. gjd_decode_batch ----------------------------------------------------------------------------------------------
for label, unstuff, pin_in in (("pageable in, host unstuffing  ", "0", False), ("pageable in, device unstuffing", "1", False),
                               ("pinned in,   device unstuffing", "1", True)):
    os.environ["GJD_DEVICE_UNSTUFF"] = unstuff
    with gjd.Decoder(local, scratch_bytes=4 << 30) as d:
        items, hold = d.make_items_pinned(files) if pin_in else d.make_items(files)
        for to_host in (False, True):
            for i in range(n):
                items[i].rgb = (pinned.ptr if to_host else dev.data_ptr()) + i * out_bytes
                items[i].rgb_on_device = 0 if to_host else 1
            d.decode_batch_raw(items, n)
            best = 1e9
            for _ in range(3):
                barrier()
                t0 = time.perf_counter()
                d.decode_batch_raw(items, n)
                best = min(best, allmax(time.perf_counter() - t0))
            k = d.stage_times()
            say("%s, pixels %s: %7.2f ms per rank = %7.1f GPixel/s total   (kernels of the last sub-batch %.2f ms)" %
                (label, "to pinned host" if to_host else "stay on device", best * 1e3, world * n * 1.048576 / best / 1e3, k[3]))
if dist:
    dist.destroy_process_group()
