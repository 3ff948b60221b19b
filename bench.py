#!/usr/bin/env python
"""bench.py — decoded MPixel/s of the 4:4:4 baseline-JPEG decode path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--images 3000]

Workload (BASELINE.json configs[2], SURVEY §8d "C3"): a batch of 3000 synthetic 1024x1024 q90 4:4:4
JPEGs per GPU (64 distinct files, seeds 0..63, tiled), encoded with the reference's vendored IJG cjpeg.
One "step" = one pass of the whole decode path over that batch.  Weak scaling: every rank (one process
per GPU, launched by torch.distributed.run for N > 1) decodes its own batch of the same size; there is
no data-path collective — torch.distributed is used for the barrier and the max-over-ranks.

  value     device-resident: parsed + unstuffed scans already in HBM, outputs stay in HBM, timed with
            CUDA events on the launching stream around every GPU stage (scratch clears included).
  e2e       gjd_decode_batch() through the C ABI with HOST buffers: header parse, unstuffing into pinned
            staging, H2D, all GPU stages, D2H of the pixels into pinned host memory, all timed.
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
WORKLOAD = "c3: synthetic 1024x1024 q90 4:4:4 baseline JPEG, batch of %d per GPU (64 distinct files tiled)"


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


def port_bench(threads, seconds, files):
    """Fallback CPU baseline: the oracle restatement (kind "port") on `threads` Python threads (ctypes drops the GIL)."""
    import ctypes as C
    so = os.path.join(ROOT, "oracle", "_build", "libgjd_oracle.so")
    if not os.path.exists(so):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, stdout=subprocess.DEVNULL)
    lib = C.CDLL(so)
    lib.gjdo_decode.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.c_void_p]
    import numpy as np
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


def cpu_reference(seconds, files, n_distinct):
    """Time the reference CPU decoder on all host cores over a bounded sample of the workload."""
    cores = os.cpu_count() or 1
    import gen_synth
    paths = [os.path.join(gen_synth.CACHE, "s%d_1024x1024_q90_o0_r0.jpg" % s) for s in range(n_distinct)]
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_bench")) and all(os.path.exists(p) for p in paths):
        r = ref_bench(cores, seconds, paths)
        kind = "reference"
    else:
        r = port_bench(cores, seconds, files)
        kind = "port"
    sample = "%d images of the c3 set (1024x1024 q90), %.1f s busy per thread, extract()+decode() timed" % (r["images"], r["busy_s_max"])
    return {"value": round(r["mpix_per_s"], 3), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}, r


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    import gen_synth
    files = gen_synth.c3_files()
    per_step = float(os.environ.get("GJD_REF_STEP_SECONDS", "3"))
    for _ in range(args.warmup):
        cpu_reference(min(per_step, 1.0), files, len(files))
    pixels, busy, images = 0, 0.0, 0
    t0 = time.time()
    for _ in range(args.steps):
        cb, r = cpu_reference(per_step, files, len(files))
        pixels += r["pixels"]
        busy += r["busy_s_max"]
        images += r["images"]
    value = pixels / busy / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(busy / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": WORKLOAD % args.images, "step": "bounded sample: every host thread decodes c3 images for %.1f s" % per_step},
        "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": cb["cores"], "kind": cb["kind"],
                         "sample": "%d images over %d steps, extract()+decode() timed per image as cpp-decoder/benchmark/benchmark.cc:31-35" % (images, args.steps)},
        "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": round(time.time() - t0, 2),
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gjd")
    ap.add_argument("--images", type=int, default=int(os.environ.get("GJD_BENCH_IMAGES", "3000")))
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import ctypes as C
    import numpy as np
    import torch
    import gen_synth
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

    base = gen_synth.c3_files()
    files = [base[i % len(base)] for i in range(args.images)]
    # the ranks of one node share its cores: each rank's host pool (header parse, unstuffing) gets its share
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    if local_world > 1 and "GJD_HOST_THREADS" not in os.environ:
        os.environ["GJD_HOST_THREADS"] = str(max(2, (os.cpu_count() or 2) // local_world))
    dec = gjd.Decoder(local, scratch_bytes=int(os.environ.get("GJD_SCRATCH", str(4 << 30))))

    # ---- device-resident throughput ("value") ------------------------------------------------------
    sampler = ClockSampler(local)     # started early: nvidia-smi needs a moment before its first sample
    batch = dec.prepare(files)
    info = None
    for _ in range(max(args.warmup, 3)):
        batch.run()
        assert batch.sync() == 0, "device reported a corrupt image"
    info = batch.info()
    stage = np.zeros(4)
    barrier()
    t0 = time.time()
    w0 = time.perf_counter()
    for _ in range(args.steps):
        batch.run()
        batch.sync()
        stage += np.array(dec.stage_times())
    torch.cuda.synchronize()
    wall = time.perf_counter() - w0
    t1 = time.time()
    barrier()
    clocks = sampler.stop(t0, t1)
    dev_ms = float(stage[3])          # CUDA-event time of all K steps on this rank
    pixels = info["pixels"]
    # one spot check that the timed path produced real pixels
    chk = batch.fetch(args.images - 1)
    assert chk.shape == (3, 1024, 1024) and int(chk.max()) > 0
    batch.free()

    # ---- end to end through the C ABI with host buffers ("e2e") -------------------------------------
    out_bytes = 3 * 1024 * 1024
    pinned = gjd.PinnedBuffer(out_bytes * args.images)
    items, hold = dec.make_items(files)
    for i in range(args.images):
        items[i].rgb = pinned.ptr + i * out_bytes
    e2e_steps = max(1, min(args.e2e_steps, args.steps))
    dec.decode_batch_raw(items, args.images)      # warm-up: grows the staging/arena buffers
    barrier()
    e0 = time.perf_counter()
    for _ in range(e2e_steps):
        dec.decode_batch_raw(items, args.images)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - e0
    barrier()
    assert all(items[i].status == 0 for i in range(0, args.images, 97))
    assert int(pinned.array[-out_bytes:].max()) > 0
    # whole-job bytes per step, like the values they sit beside (every rank copies its own shard)
    h2d = sum(len(f) for f in files) * world   # upper bound of the unstuffed scans + tables actually copied
    d2h = out_bytes * args.images * world

    # ---- the same call with the pixels left in HBM (rgb_on_device: the data-loader use; extra key) ---
    dev_out = torch.zeros(out_bytes * args.images, dtype=torch.uint8, device="cuda")
    for i in range(args.images):
        items[i].rgb = dev_out.data_ptr() + i * out_bytes
        items[i].rgb_on_device = 1
    dec.decode_batch_raw(items, args.images)
    barrier()
    e0 = time.perf_counter()
    for _ in range(e2e_steps):
        dec.decode_batch_raw(items, args.images)
    torch.cuda.synchronize()
    dout_s = time.perf_counter() - e0
    barrier()
    assert int(dev_out[-out_bytes:].max()) > 0
    del dev_out

    # ---- reduce over ranks ---------------------------------------------------------------------------
    t = torch.tensor([dev_ms, e2e_s, wall, dout_s], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_s_max, wall_max, dout_s_max = (float(x) for x in t.cpu())
    total_pixels = pixels * world
    value = total_pixels * args.steps / (dev_ms_max * 1e-3) / 1e6
    e2e_value = total_pixels * e2e_steps / e2e_s_max / 1e6

    if rank == 0:
        peak, peak_src = measured_peak()
        t_tr = stage[2] / args.steps * 1e-3
        algo = info["padded_pixels"] * 6 + info["pixels"] * 3
        achieved = algo / t_tr / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "transform_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(dev_ms_max / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": WORKLOAD % args.images, "images_per_gpu": args.images,
                       "scan_bytes_per_gpu": info["scan_bytes"], "subsequences_per_gpu": info["subseqs"],
                       "l2": "inputs+outputs larger than L2 (scans %.2f GB read 3x, RGB %.1f GB written per step)" % (info["scan_bytes"] / 1e9, info["pixels"] * 3 / 1e9),
                       "timing": "CUDA events on the launching stream per step (value); wall clock around gjd_decode_batch (e2e, %d steps)" % e2e_steps,
                       "parallelism": "replicas: one sub-batch per GPU, no collective"},
            "stages_ms": {"entropy_sync": round(stage[0] / args.steps, 3), "entropy_write": round(stage[1] / args.steps, 3),
                          "transform": round(stage[2] / args.steps, 3), "pipeline": round(stage[3] / args.steps, 3),
                          "wall_per_step": round(wall_max / args.steps * 1e3, 3)},
            "roofline": {"kernel": "k_transform", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo},
            "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": round(e2e_s_max / e2e_steps * 1e3, 2)},
            "e2e_device_output": {"value": round(total_pixels * e2e_steps / dout_s_max / 1e6, 1), "unit": UNIT,
                                  "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 0, "steps": e2e_steps,
                                  "ms_per_step": round(dout_s_max / e2e_steps * 1e3, 2),
                                  "note": "gjd_decode_batch with host JPEG bytes in, pixels left in device memory (no read-back)"},
            "gpu_launches": int(info["launches"]) * args.steps,
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            cb, _ = cpu_reference(float(os.environ.get("GJD_CPU_BASELINE_SECONDS", "12")), base, len(base))
            line["cpu_baseline"] = cb
            paths = [os.path.join(gen_synth.CACHE, "s%d_1024x1024_q90_o0_r0.jpg" % s) for s in range(len(base))]
            if all(os.path.exists(p) for p in paths):
                jl = jpeglib_bench(cb["cores"], float(os.environ.get("GJD_JPEGLIB_BASELINE_SECONDS", "6")), paths)
                if jl:
                    line["cpu_baseline_jpeglib"] = {
                        "value": round(jl["mpix_per_s"], 3), "unit": UNIT, "cores": cb["cores"], "kind": "reference-vendored libjpeg 9e",
                        "sample": "%d images of the c3 set, %.1f s busy per thread, jpeg_read_header..jpeg_finish_decompress timed (RGB out, JDCT_ISLOW)" % (jl["images"], jl["busy_s_max"])}
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    pinned.free()
    dec.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
