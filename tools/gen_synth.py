"""Deterministic synthetic 4:4:4 baseline JPEGs.

Images: multi-octave value noise + N(0,4) pixel noise (natural-image-like spectrum, ~0.25-0.4 B/px at
q90), written as P6 and encoded with the IJG libjpeg 9e `cjpeg` binary the reference vendors
(jpeglib-implementation/libjpeg/bin/cjpeg), exactly as `cjpeg -quality Q -sample 1x1 -baseline`.
build() copies that binary (not source) to tools/_build/libjpeg so it travels to the GPU box.
Encoded files are cached under tools/_build/synth/.
"""
import hashlib
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBJPEG = os.path.join(ROOT, "tools", "_build", "libjpeg")
CACHE = os.path.join(ROOT, "tools", "_build", "synth")


def have_cjpeg():
    return os.path.exists(os.path.join(LIBJPEG, "bin", "cjpeg"))


def make_image(seed, w, h):
    rng = np.random.default_rng(seed)
    acc = np.zeros((h, w, 3), np.float32)
    amp_total = 0.0
    for o in range(6):
        gh, gw = (h >> (7 - o)) + 2, (w >> (7 - o)) + 2
        g = rng.random((gh, gw, 3), dtype=np.float32)
        ys = np.linspace(0, gh - 1.001, h, dtype=np.float32)
        xs = np.linspace(0, gw - 1.001, w, dtype=np.float32)
        y0 = ys.astype(np.int32)
        x0 = xs.astype(np.int32)
        fy = (ys - y0)[:, None, None]
        fx = (xs - x0)[None, :, None]
        a = g[y0][:, x0]
        b = g[y0][:, x0 + 1]
        c = g[y0 + 1][:, x0]
        d = g[y0 + 1][:, x0 + 1]
        amp = 0.5 ** o
        acc += amp * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)
        amp_total += amp
    img = acc / amp_total * 255.0
    img += rng.normal(0, 4, img.shape).astype(np.float32)
    return np.clip(img, 0, 255).astype(np.uint8)


def encode(img, quality=90, optimize=False, restart=0, grayscale=False, sample="1x1"):
    """RGB uint8 (H,W,3) -> JPEG bytes via vendored cjpeg (sample: the luma sampling factors, "2x2" = 4:2:0, "2x1" = 4:2:2)."""
    h, w, _ = img.shape
    ppm = b"P6\n%d %d\n255\n" % (w, h) + img.tobytes()
    cmd = [os.path.join(LIBJPEG, "bin", "cjpeg"), "-quality", str(quality), "-sample", sample, "-baseline"]
    if optimize:
        cmd.append("-optimize")
    if restart:
        cmd += ["-restart", str(restart)]
    if grayscale:
        cmd.append("-grayscale")
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(LIBJPEG, "lib"))
    return subprocess.run(cmd, input=ppm, stdout=subprocess.PIPE, check=True, env=env).stdout


def synth_jpeg(seed, w, h, quality=90, optimize=False, restart=0, sample="1x1"):
    key = f"s{seed}_{w}x{h}_q{quality}_o{int(optimize)}_r{restart}" + ("" if sample == "1x1" else "_" + sample)
    path = os.path.join(CACHE, key + ".jpg")
    if os.path.exists(path):
        with open(path, "rb") as f:
            return f.read()
    data = encode(make_image(seed, w, h), quality, optimize, restart, False, sample)
    os.makedirs(CACHE, exist_ok=True)
    tmp = path + ".%d.tmp" % os.getpid()
    with open(tmp, "wb") as f:
        f.write(data)
    os.replace(tmp, path)
    return data


def c3_files(n_distinct=64, size=1024, quality=90):
    """BASELINE config 3: 1024x1024 q90, seeds 0..63."""
    return [synth_jpeg(s, size, size, quality) for s in range(n_distinct)]


def c5_specs(n_distinct=256):
    """BASELINE config 5: mixed sizes, q in {75,90,95}, default_rng(2024)."""
    rng = np.random.default_rng(2024)
    sizes = [256, 384, 512, 640, 768, 1024, 1536, 2048]
    specs = []
    for i in range(n_distinct):
        q = int(rng.choice([75, 90, 95]))
        if rng.random() < 0.10:
            w, h = int(rng.integers(200, 1101)), int(rng.integers(200, 1101))
        else:
            w = h = int(rng.choice(sizes))
        specs.append((2000 + i, w, h, q))
    return specs


def c5_files(n_distinct=256):
    """The distinct files of config 5 (missing ones are encoded on a thread pool: cjpeg runs as a subprocess)."""
    from concurrent.futures import ThreadPoolExecutor
    specs = c5_specs(n_distinct)
    with ThreadPoolExecutor(max(1, min(16, os.cpu_count() or 1))) as ex:
        return list(ex.map(lambda s: synth_jpeg(*s), specs))


def digest(files):
    m = hashlib.sha256()
    for f in files:
        m.update(f)
    return m.hexdigest()[:16]
