#!/usr/bin/env python
"""bench.py — decoded MPixel/s of the 4:4:4 baseline-JPEG decode path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config c3|c5] [--images M]

Workloads:
  c3 (default; BASELINE.json configs[2], the shape `metric` is quoted on): a batch of 3000 synthetic
      1024x1024 q90 4:4:4 JPEGs PER GPU (64 distinct files, seeds 0..63, tiled).  Weak scaling: every rank
      decodes its own batch of the same size.
  c5 (BASELINE.json configs[4]): ONE global list of 24000 mixed-size images (256 distinct files,
      default_rng(2024): sizes {256..2048}^2 + 10 % ragged, q in {75,90,95}), sharded over the N ranks with
      gjd_partition (greedy LPT on padded pixels + 5 x file bytes, include/gjd.h); rank r decodes part_of == r.  Strong scaling.
Both are encoded with the reference's vendored IJG cjpeg.  One "step" = one pass of the whole decode path
over the rank's images.  There is no data-path collective — torch.distributed carries the
barrier, the max-over-ranks of the times and the sum of the parity counts.

  value     device-resident: parsed + unstuffed scans already in HBM, outputs stay in HBM, timed with
            CUDA events on the launching stream around every GPU stage (scratch clears included).  A shard
            that does not fit in HBM at once (c5 at small N) is decoded as successive resident chunks; a
            chunk's preparation (the reference's untimed extract()/allocate() loop,
            cuda-decoder/benchmark_thoughput/benchmark.cu:49-60) is outside the event-timed region.  With N > 1
            the ranks pass a barrier before the timed runs of every chunk, so that no rank's untimed host work
            (parity checks, freeing and preparing chunks) runs beside another rank's timed launches
            (GJD_BENCH_LOCKSTEP=0 turns that off, GJD_BENCH_TRACE=1 prints every run's stage times per rank).
  e2e       gjd_decode_batch() through the C ABI with HOST buffers: JPEG files in page-locked host memory in, header
            parse, H2D of the raw scans, unstuffing + all GPU stages, D2H of the pixels into page-locked host memory,
            all timed (wall clock around the calls).
  e2e_device_output
            the same call with the pixels left in device memory (rgb_on_device).
  parity_checked
            after EACH of the three timed loops, one image per distinct file (spread over the batch) is
            compared byte for byte with the oracle's output for that file (oracle/gjd_oracle.c, hashed once
            outside every timed region); a mismatch fails the run.
  roofline  the transform kernel (dequant+IDCT+colour+store), algorithmic bytes 6 B/padded px + 3 B/px,
            against the measured HBM copy peak of MEASURED_PEAKS.json.
  cpu_baseline / --impl reference
            the UNMODIFIED reference cpp-decoder (oracle/_ref/ref_bench: JPEGParser::extract+decode, the
            timed region of cpp-decoder/benchmark/benchmark.cc:31-35) on all host cores, bounded sample.
  cpu_baseline_jpeglib
            the IJG libjpeg 9e the reference vendors, in process on all host cores (oracle/_ref/jpeglib_bench),
            a second, reported-only baseline.
"""
import argparse
import hashlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gpu-jpeg-decoder_b200")
for p in (PKG, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "decoded MPixel/s (4:4:4 batch)"
UNIT = "MPixel/s"
CHUNK_PIXELS = int(float(os.environ.get("GJD_BENCH_CHUNK_GPIX", "3.3")) * 1e9)   # padded pixels resident at once


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for r in self.rows if t0 <= r[0] <= t1 + 0.2]
        window = "timed region"
        if not rows:   # a timed region shorter than the sampling period: the samples next to it
            rows = [r for r in self.rows if t0 - 0.5 <= r[0] <= t1 + 0.5]
            window = "timed region +-0.5 s"
        for ts, line in rows:
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                for n, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "window": window}


# ---------------------------------------------------------------------------------------- workloads ----

def load_workload(config, images):
    """(distinct files as bytes, their paths on disk, file id of every image of the GLOBAL list)."""
    import gen_synth
    if config == "c3":
        base = gen_synth.c3_files()
        paths = [os.path.join(gen_synth.CACHE, "s%d_1024x1024_q90_o0_r0.jpg" % s) for s in range(len(base))]
    elif config == "c5":
        specs = gen_synth.c5_specs(256)
        base = gen_synth.c5_files(256)
        paths = [os.path.join(gen_synth.CACHE, "s%d_%dx%d_q%d_o0_r0.jpg" % s) for s in specs]
    else:
        raise SystemExit("bench.py: unknown --config %r" % config)
    return base, paths, [i % len(base) for i in range(images)]


def config_dict(config, images, world):
    """The static description of the workload — identical in the gjd arm and in the --impl reference arm."""
    if config == "c3":
        return {"workload": "c3: synthetic 1024x1024 q90 4:4:4 baseline JPEG, batch of %d per GPU (64 distinct files tiled)" % images,
                "config": "c3", "images_per_gpu": images, "distinct_files": 64, "scaling": "weak",
                "parallelism": "replicas: one batch per GPU, no collective",
                "l2": "inputs+outputs larger than L2 (0.75 GB of scans read 3x, 9.4 GB of RGB written per step and GPU)"}
    return {"workload": "c5: synthetic mixed-size 4:4:4 baseline JPEG ({256..2048}^2 + 10%% ragged, q 75/90/95), %d images "
                        "(256 distinct files tiled) sharded over the GPUs by gjd_partition" % images,
            "config": "c5", "images_total": images, "distinct_files": 256, "scaling": "strong",
            "parallelism": "shards: gjd_partition (greedy LPT on padded pixels + 5 x file bytes), one shard per GPU, no collective",
            "l2": "inputs+outputs larger than L2 (every resident chunk holds > 1 GPixel)"}


def default_images(config):
    return int(os.environ.get("GJD_BENCH_IMAGES", "3000" if config == "c3" else "24000"))


# ---------------------------------------------------------------------------------------- CPU arms -----

def ref_bench(threads, seconds, files_on_disk):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
    out = subprocess.run([exe, str(threads), str(seconds), *files_on_disk], stdout=subprocess.PIPE, check=True, text=True)
    return json.loads(out.stdout.strip().splitlines()[-1])


def jpeglib_bench(threads, seconds, files_on_disk):
    """Second CPU baseline: the IJG libjpeg 9e the reference vendors, in process, all host cores."""
    exe = os.path.join(ROOT, "oracle", "_ref", "jpeglib_bench")
    if not os.path.exists(exe):
        return None
    out = subprocess.run([exe, str(threads), str(seconds), *files_on_disk], stdout=subprocess.PIPE, check=True, text=True)
    return json.loads(out.stdout.strip().splitlines()[-1])


def oracle_lib():
    import ctypes as C
    so = os.path.join(ROOT, "oracle", "_build", "libgjd_oracle.so")
    if not os.path.exists(so):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, stdout=subprocess.DEVNULL)
    lib = C.CDLL(so)
    lib.gjdo_decode.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def port_bench(threads, seconds, files):
    """Fallback CPU baseline: the oracle restatement (kind "port") on `threads` Python threads (ctypes drops the GIL)."""
    import ctypes as C
    import numpy as np
    lib = oracle_lib()
    busy, pixels, images = [0.0] * threads, [0] * threads, [0] * threads

    def work(t):
        k = t
        rgb = np.empty(3 * 4096 * 4096, np.uint8)
        while busy[t] < seconds:
            d = files[k % len(files)]
            k += threads
            w, h = C.c_int(), C.c_int()
            t0 = time.perf_counter()
            lib.gjdo_decode(C.cast(C.c_char_p(d), C.c_void_p), len(d), C.byref(w), C.byref(h), None, None, rgb.ctypes.data)
            busy[t] += time.perf_counter() - t0
            pixels[t] += w.value * h.value
            images[t] += 1

    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [t.start() for t in th]
    [t.join() for t in th]
    return {"threads": threads, "images": sum(images), "pixels": sum(pixels), "busy_s_max": max(busy),
            "mpix_per_s": sum(pixels) / max(busy) / 1e6}


def cpu_reference(seconds, base, paths, config):
    """Time the reference CPU decoder on all host cores over a bounded sample of the workload."""
    cores = os.cpu_count() or 1
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_bench")) and all(os.path.exists(p) for p in paths):
        r = ref_bench(cores, seconds, paths)
        kind = "reference"
    else:
        r = port_bench(cores, seconds, base)
        kind = "port"
    sample = "%d images of the %s set, %.1f s busy per thread, extract()+decode() timed" % (r["images"], config, r["busy_s_max"])
    return {"value": round(r["mpix_per_s"], 3), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}, r


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    base, paths, _ = load_workload(args.config, 1)
    per_step = float(os.environ.get("GJD_REF_STEP_SECONDS", "3"))
    for _ in range(args.warmup):
        cpu_reference(min(per_step, 1.0), base, paths, args.config)
    pixels, busy, images = 0, 0.0, 0
    t0 = time.time()
    for _ in range(args.steps):
        cb, r = cpu_reference(per_step, base, paths, args.config)
        pixels += r["pixels"]
        busy += r["busy_s_max"]
        images += r["images"]
    value = pixels / busy / 1e6
    cfg = config_dict(args.config, args.images, world)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(busy / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": cb["cores"], "kind": cb["kind"],
                         "sample": "bounded sample: every host thread decodes %s images for %.1f s per step; %d images over %d steps, "
                                   "extract()+decode() timed per image as cpp-decoder/benchmark/benchmark.cc:31-35"
                                   % (args.config, per_step, images, args.steps)},
        "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "host_cores": cb["cores"], "wall_s": round(time.time() - t0, 2),
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------- parity -------

def oracle_digests(base, ids, threads):
    """md5 of the oracle's planar R,G,B output (oracle/gjd_oracle.c) for the distinct files in `ids`."""
    import ctypes as C
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    lib = oracle_lib()

    def one(i):
        d = base[i]
        w, h = C.c_int(), C.c_int()
        buf = C.cast(C.c_char_p(d), C.c_void_p)
        rc = lib.gjdo_decode(buf, len(d), C.byref(w), C.byref(h), None, None, None)
        assert rc == 0, "oracle rejected file %d" % i
        rgb = np.empty(3 * w.value * h.value, np.uint8)
        rc = lib.gjdo_decode(buf, len(d), C.byref(w), C.byref(h), None, None, rgb.ctypes.data)
        assert rc == 0
        return i, hashlib.md5(rgb.tobytes()).digest()

    with ThreadPoolExecutor(max(1, threads)) as ex:
        return dict(ex.map(one, sorted(set(ids))))


def pick_samples(ids):
    """One position per distinct file id present in `ids`, spread over the list."""
    occ = {}
    for pos, fid in enumerate(ids):
        occ.setdefault(fid, []).append(pos)
    return sorted(v[(fid * 7 + 3) % len(v)] for fid, v in occ.items())


def make_chunks(ids, weight_of):
    """Split the rank's image list into resident chunks of at most CHUNK_PIXELS padded pixels."""
    chunks, cur, acc = [], [], 0
    for pos, fid in enumerate(ids):
        w = weight_of[fid]
        if cur and acc + w > CHUNK_PIXELS:
            chunks.append(cur)
            cur, acc = [], 0
        cur.append(pos)
        acc += w
    if cur:
        chunks.append(cur)
    return chunks


# ---------------------------------------------------------------------------------------- main ---------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gjd")
    ap.add_argument("--config", default=os.environ.get("GJD_BENCH_CONFIG", "c3"), choices=["c3", "c5"])
    ap.add_argument("--images", type=int, default=0, help="c3: images per GPU (3000); c5: images in the global list (24000)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.images <= 0:
        args.images = default_images(args.config)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import numpy as np
    import torch
    import gjd

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; gjd has no CPU path (use --impl reference for the CPU arm)")
    # stdout carries the ONE JSON line and nothing else: libraries that write to fd 1 (NCCL prints its version
    # there when NCCL_DEBUG is set) are sent to stderr for the length of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- the rank's images --------------------------------------------------------------------------
    base, paths, gids = load_workload(args.config, args.images)
    shards = None
    if args.config == "c5":
        files = [base[g] for g in gids]
        # every rank computes the same partition: greedy LPT on padded pixels + a term for the scan bytes
        part_of, load, cost = gjd.partition(files, world, gjd.PARTITION_BYTE_WEIGHT)
        ids = [gids[i] for i in np.nonzero(part_of == rank)[0]]
        nbytes = np.array([len(f) for f in files], np.uint64)
        pixels = np.where(cost > 0, cost - np.uint64(gjd.PARTITION_BYTE_WEIGHT) * nbytes, 0).astype(np.uint64)
        spread = lambda v: round(float(max(v) / (sum(v) / world) - 1.0), 6)
        per_px = [int(pixels[part_of == r].sum()) for r in range(world)]
        per_by = [int(nbytes[part_of == r].sum()) for r in range(world)]
        shards = {"weight": "padded pixels + %d x file bytes" % gjd.PARTITION_BYTE_WEIGHT,
                  "padded_pixels_per_rank": per_px, "file_bytes_per_rank": per_by,
                  "imbalance": spread([int(x) for x in load]),
                  "imbalance_pixels": spread(per_px), "imbalance_bytes": spread(per_by)}
    else:
        ids = gids
    n_local = len(ids)
    dims = {}
    for fid in set(ids):
        rc, w, h = gjd.peek(base[fid])
        assert rc == 0
        dims[fid] = (w, h)
    weight_of = {fid: (8 * ((w + 7) // 8)) * (8 * ((h + 7) // 8)) for fid, (w, h) in dims.items()}
    out_bytes = {fid: 3 * w * h for fid, (w, h) in dims.items()}
    chunks = make_chunks(ids, weight_of)
    samples = set(pick_samples(ids))
    # the ranks of one node share its cores: each rank's host pool (header parse, unstuffing) gets its share
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    host_threads = max(2, (os.cpu_count() or 2) // local_world)
    if local_world > 1 and "GJD_HOST_THREADS" not in os.environ:
        os.environ["GJD_HOST_THREADS"] = str(host_threads)
    want = oracle_digests(base, [ids[p] for p in samples], host_threads)      # the checker, outside every timed region
    dec = gjd.Decoder(local, scratch_bytes=int(os.environ.get("GJD_SCRATCH", str(4 << 30))))

    def check(pos, pixels, leg):
        got = hashlib.md5(np.ascontiguousarray(pixels).tobytes()).digest()
        if got != want[ids[pos]]:
            raise SystemExit("bench.py: PARITY FAILURE in the %s leg: image %d (file %d) differs from the oracle" % (leg, pos, ids[pos]))

    # ---- device-resident throughput ("value") ------------------------------------------------------
    sampler = ClockSampler(local)     # started early: nvidia-smi needs a moment before its first sample
    stage = np.zeros(4)
    info_sum = dict(pixels=0, padded_pixels=0, scan_bytes=0, subseqs=0, launches=0)
    checked = {"value": 0, "e2e": 0, "e2e_device_output": 0}
    wall = 0.0
    barrier()
    t0 = time.time()
    trace = os.environ.get("GJD_BENCH_TRACE") == "1"
    lockstep = os.environ.get("GJD_BENCH_LOCKSTEP", "1") == "1" and dist is not None
    rounds = len(chunks)
    if lockstep:      # ranks may hold different numbers of chunks: all of them make the same number of barrier calls
        r_ = torch.tensor([rounds], device="cuda")
        dist.all_reduce(r_, op=dist.ReduceOp.MAX)
        rounds = int(r_.item())
    for ci in range(rounds):
        if ci >= len(chunks):
            barrier()
            continue
        chunk = chunks[ci]
        batch = dec.prepare([base[ids[p]] for p in chunk])
        for _ in range(max(args.warmup, 3) if ci == 0 else 1):
            batch.run()
            assert batch.sync() == 0, "device reported a corrupt image"
        if ci == 0 or lockstep:
            # every rank times chunk ci at the same time: the host work between chunks (parity checks against the
            # oracle, header parsing of the next chunk) of one rank otherwise runs beside another rank's timed launches
            barrier()
        if ci == 0:
            t0 = time.time()
        w0 = time.perf_counter()
        for _ in range(args.steps):
            batch.run()
            batch.sync()
            st = np.array(dec.stage_times())
            stage += st
            if trace:
                print("trace rank %d chunk %d: before_sync %.3f sync %.3f write %.3f transform %.3f pipeline %.3f ms"
                      % (rank, ci, st[3] - st[0] - st[1] - st[2], st[0], st[1], st[2], st[3]), file=sys.stderr, flush=True)
        torch.cuda.synchronize()
        wall += time.perf_counter() - w0
        info = batch.info()
        for k in info_sum:
            info_sum[k] += info[k]
        for j, p in enumerate(chunk):       # parity of what was just timed: the batch's own output buffers
            if p in samples:
                assert batch.status(j) == 0
                check(p, batch.fetch(j), "value")
                checked["value"] += 1
        batch.free()
    t1 = time.time()
    barrier()
    clocks = sampler.stop(t0, t1)
    dev_ms = float(stage[3])          # CUDA-event time of all K steps over all chunks on this rank
    pixels = info_sum["pixels"]

    # ---- end to end through the C ABI with host buffers ("e2e") -------------------------------------
    chunk_out = [sum(out_bytes[ids[p]] for p in c) for c in chunks]
    pinned = gjd.PinnedBuffer(max(chunk_out))
    citems = []
    for c in chunks:
        # the step's inputs lie in page-locked host memory (gjd_host_alloc), files back to back
        items, hold = dec.make_items_pinned([base[ids[p]] for p in c])
        off, offs = 0, []
        for j, p in enumerate(c):
            items[j].rgb = pinned.ptr + off
            offs.append(off)
            off += out_bytes[ids[p]]
        citems.append((items, hold, offs))
    e2e_steps = max(1, min(args.e2e_steps, args.steps))

    def e2e_loop(leg, fetch):
        for items, hold, offs in citems:      # warm-up: grows the staging/arena buffers
            dec.decode_batch_raw(items, len(offs))
        barrier()
        total = 0.0
        for step in range(e2e_steps):
            for c, (items, hold, offs) in zip(chunks, citems):
                e0 = time.perf_counter()
                dec.decode_batch_raw(items, len(offs))
                torch.cuda.synchronize()
                total += time.perf_counter() - e0
                if step == e2e_steps - 1:     # parity of the last timed step (clock stopped)
                    for j, p in enumerate(c):
                        if p in samples:
                            assert items[j].status == 0
                            check(p, fetch(offs[j], out_bytes[ids[p]]), leg)
                            checked[leg] += 1
                    assert all(items[j].status == 0 for j in range(0, len(offs), 97))
        barrier()
        return total

    e2e_s = e2e_loop("e2e", lambda off, n: pinned.array[off:off + n])
    # whole-job bytes per step, like the values they sit beside
    h2d_local = sum(len(base[f]) for f in ids)   # upper bound of the unstuffed scans + tables actually copied
    d2h_local = sum(chunk_out)

    # ---- the same call with the pixels left in HBM (rgb_on_device: the data-loader use; extra key) ---
    dev_out = torch.zeros(max(chunk_out), dtype=torch.uint8, device="cuda")
    for items, hold, offs in citems:
        for j, off in enumerate(offs):
            items[j].rgb = dev_out.data_ptr() + off
            items[j].rgb_on_device = 1
    dout_s = e2e_loop("e2e_device_output", lambda off, n: dev_out[off:off + n].cpu().numpy())
    del dev_out

    # ---- reduce over ranks ---------------------------------------------------------------------------
    t = torch.tensor([dev_ms, e2e_s, wall, dout_s], dtype=torch.float64, device="cuda")
    s = torch.tensor([pixels, h2d_local, d2h_local, checked["value"], checked["e2e"], checked["e2e_device_output"],
                      info_sum["launches"] * args.steps, n_local, info_sum["scan_bytes"], info_sum["subseqs"]],
                     dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_s_max, wall_max, dout_s_max = (float(x) for x in t.cpu())
    # what each rank saw (a shard that lags shows here, not in the maxima): ms per step and stage, scan bytes, chunks, clock
    mine = torch.tensor([dev_ms / args.steps, stage[0] / args.steps, stage[1] / args.steps, stage[2] / args.steps,
                         info_sum["scan_bytes"], len(chunks), (clocks or {}).get("sm_mhz") or 0.0, e2e_s / e2e_steps * 1e3],
                        dtype=torch.float64, device="cuda")
    per_rank = [torch.zeros_like(mine) for _ in range(world)]
    if dist:
        dist.all_gather(per_rank, mine)
    else:
        per_rank = [mine]
    per_rank = [[round(float(x), 3) for x in r.cpu()] for r in per_rank]
    total_pixels, h2d, d2h, chk_v, chk_e, chk_d, launches, n_total, scan_total, subs_total = (int(x) for x in s.cpu())
    value = total_pixels * args.steps / (dev_ms_max * 1e-3) / 1e6
    e2e_value = total_pixels * e2e_steps / e2e_s_max / 1e6

    if rank == 0:
        peak, peak_src = measured_peak()
        t_tr = stage[2] / args.steps * 1e-3
        algo = info_sum["padded_pixels"] * 6 + info_sum["pixels"] * 3
        achieved = algo / t_tr / 1e9
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "transform_traffic.json")) as f:
                tj = json.load(f)
            if args.config == "c3" and args.images == 3000:
                traffic, traffic_src = tj.get("dram_bytes_per_launch"), tj.get("source")
        except Exception:
            pass
        cfg = config_dict(args.config, args.images, world)
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(dev_ms_max / args.steps, 3), "higher_is_better": True,
            "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": cfg,
            "workload_detail": {"images_decoded_per_step": n_total, "pixels_per_step": total_pixels, "scan_bytes": scan_total,
                                "subsequences": subs_total, "resident_chunks_rank0": len(chunks),
                                "timing": "CUDA events on the launching stream per step (value); wall clock around gjd_decode_batch (e2e, %d steps)" % e2e_steps},
            "parity_checked": min(chk_v, chk_e, chk_d),
            "parity_detail": {"value": chk_v, "e2e": chk_e, "e2e_device_output": chk_d,
                              "how": "one image per distinct file and rank, spread over the batch, md5 of the planar pixels vs the oracle's (oracle/gjd_oracle.c) after each timed loop"},
            "stages_ms": {"entropy_sync": round(stage[0] / args.steps, 3), "entropy_write": round(stage[1] / args.steps, 3),
                          "transform": round(stage[2] / args.steps, 3), "pipeline": round(stage[3] / args.steps, 3),
                          "wall_per_step": round(wall_max / args.steps * 1e3, 3)},
            "roofline": {"kernel": "k_transform", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo, "launches_per_step_rank0": len(chunks)},
            "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": round(e2e_s_max / e2e_steps * 1e3, 2)},
            "e2e_device_output": {"value": round(total_pixels * e2e_steps / dout_s_max / 1e6, 1), "unit": UNIT,
                                  "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 0, "steps": e2e_steps,
                                  "ms_per_step": round(dout_s_max / e2e_steps * 1e3, 2),
                                  "note": "gjd_decode_batch with host JPEG bytes in, pixels left in device memory (no read-back)"},
            "gpu_launches": launches, "host_cores": os.cpu_count(), "host_threads_per_rank": int(os.environ.get("GJD_HOST_THREADS", os.cpu_count() or 1)),
            "clocks": clocks,
        }
        if world > 1:
            line["per_rank"] = {"columns": ["ms_per_step", "entropy_sync_ms", "entropy_write_ms", "transform_ms", "scan_bytes",
                                            "resident_chunks", "sm_mhz", "e2e_ms_per_step"], "rows": per_rank}
        if shards is not None:
            line["shards"] = shards
        if world == 1 and not args.no_cpu_baseline:
            cb, _ = cpu_reference(float(os.environ.get("GJD_CPU_BASELINE_SECONDS", "12")), base, paths, args.config)
            line["cpu_baseline"] = cb
            line["e2e_vs_cpu_baseline"] = round(e2e_value / cb["value"], 2)       # the like-for-like ratio (host bytes in, host pixels out)
            line["value_vs_cpu_baseline_kernels_only"] = round(value / cb["value"], 1)
            if all(os.path.exists(p) for p in paths):
                jl = jpeglib_bench(cb["cores"], float(os.environ.get("GJD_JPEGLIB_BASELINE_SECONDS", "6")), paths)
                if jl:
                    line["cpu_baseline_jpeglib"] = {
                        "value": round(jl["mpix_per_s"], 3), "unit": UNIT, "cores": cb["cores"], "kind": "reference-vendored libjpeg 9e",
                        "sample": "%d images of the %s set, %.1f s busy per thread, jpeg_read_header..jpeg_finish_decompress timed (RGB out, JDCT_ISLOW)" % (jl["images"], args.config, jl["busy_s_max"])}
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    pinned.free()
    dec.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
