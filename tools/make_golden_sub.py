"""Generate tests/golden/sub/*.jpg + tests/golden/ref_outputs_sub.npz (run in the build container).

Chroma-subsampled files (4:2:2, 4:4:0, 4:2:0) are outside the reference's decoders (cpp-decoder/src/parser.cpp:52-60
ignores the sampling factors and garbles them; data_preprocessing/image_converter.py:14-19 exists to avoid them), so
no reference output exists.  The pin is composed from the reference's own pieces:
  coefficients  jpeg_read_coefficients of the reference's vendored IJG libjpeg 9e (oracle/_ref/jpeg_coefs), re-ordered
                into scan order: line = mcu * (hs*vs + 2) + block-in-MCU, zigzag order inside a block, absolute DC;
  pixels        those coefficients x the file's quantisation tables through the oracle's IDCT (gjdo_idct_block = the
                reference's idct.cpp), each chroma sample replicated over the hs x vs pixels it covers, then the oracle's
                colour function (gjdo_color = the reference's color.cpp) and the crop to W x H.
Sample replication is what libjpeg calls non-fancy upsampling (`djpeg -nosmooth`); libjpeg's default "fancy" triangle
filter is NOT what is pinned.  Against `djpeg -nosmooth` the composed pixels differ by at most a few levels (different
IDCT and colour arithmetic: the same envelope the reference itself has against libjpeg) — printed per file.
"<name>/coef" int16, "<name>/rgb" (3,H,W) uint8.  Also copies the reference's own subsampled test picture
(testing/image.jpeg, a 4:2:0 file none of its decoders can read) as sub/ref_testing_image.jpg.
"""
import ctypes as C
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth  # noqa: E402

CASES = [  # (seed, w, h, quality, sampling, optimize, cjpeg -restart argument or 0)
    (51, 64, 48, 90, "2x2", False, 0),
    (52, 401, 363, 75, "2x2", True, 0),        # ragged edges in both directions, optimised tables
    (53, 227, 130, 90, "2x1", False, 0),       # 4:2:2, ragged
    (54, 130, 227, 85, "1x2", True, 0),        # 4:4:0
    (55, 640, 360, 95, "2x2", False, "1"),     # restart every MCU row
    (56, 333, 200, 80, "2x1", False, "7B"),    # restart interval of 7 MCUs: spans tiles and rows
    (57, 17, 9, 90, "2x2", False, 0),          # smaller than two MCUs
    (58, 1024, 512, 50, "2x2", False, 0),      # many tiles, low quality
]


def encode(img, quality, sample, optimize, restart):
    h, w, _ = img.shape
    ppm = b"P6\n%d %d\n255\n" % (w, h) + img.tobytes()
    cmd = [os.path.join(gen_synth.LIBJPEG, "bin", "cjpeg"), "-quality", str(quality), "-sample", sample, "-baseline"]
    if optimize:
        cmd.append("-optimize")
    if restart:
        cmd += ["-restart", str(restart)]
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(gen_synth.LIBJPEG, "lib"))
    return subprocess.run(cmd, input=ppm, stdout=subprocess.PIPE, check=True, env=env).stdout


def read_coefs(data):
    """jpeg_read_coefficients via oracle/_ref/jpeg_coefs -> (W, H, [(h, v, q_nat[64], blocks[rows, cols, 64] natural)])."""
    exe = os.path.join(ROOT, "oracle", "_ref", "jpeg_coefs")
    with tempfile.TemporaryDirectory() as td:
        a, b = os.path.join(td, "a.jpg"), os.path.join(td, "a.bin")
        open(a, "wb").write(data)
        subprocess.run([exe, a, b], check=True)
        raw = open(b, "rb").read()
    W, H, nc = struct.unpack_from("<3i", raw, 0)
    off, comps = 12, []
    for _ in range(nc):
        h, v, cols, rows = struct.unpack_from("<4i", raw, off)
        off += 16
        q = np.frombuffer(raw, np.uint16, 64, off).astype(np.int32)
        off += 128
        blk = np.frombuffer(raw, np.int16, rows * cols * 64, off).reshape(rows, cols, 64)
        off += rows * cols * 128
        comps.append((h, v, q, blk))
    return W, H, comps


def compose(data, orc, zz_of_nat):
    W, H, comps = read_coefs(data)
    assert len(comps) == 3 and comps[1][:2] == (1, 1) and comps[2][:2] == (1, 1)
    hs, vs = comps[0][0], comps[0][1]
    my, mx = comps[1][3].shape[0], comps[1][3].shape[1]        # MCU grid = the chroma block grid
    nblk = hs * vs + 2
    coef = np.zeros((my * mx, nblk, 64), np.int16)
    planes = []
    for ci, (h, v, q, blk) in enumerate(comps):
        rows, cols, _ = blk.shape
        # pixels of this component: IDCT of every block (the oracle's, in place on int[64])
        deq = np.ascontiguousarray(blk.astype(np.int32) * q[None, None, :])
        for r in range(rows):
            for c in range(cols):
                orc.gjdo_idct_block(deq[r, c].ctypes.data)
        planes.append(deq.reshape(rows, cols, 8, 8).transpose(0, 2, 1, 3).reshape(rows * 8, cols * 8))
        # scan order: luma blocks row-major inside the MCU, then Cb, then Cr; zigzag inside a block
        zz = np.zeros_like(blk)
        zz[:, :, zz_of_nat] = blk
        for r in range(rows):
            for c in range(cols):
                mcu = (r // v) * mx + (c // h)
                bp = (r % v) * h + (c % h) if ci == 0 else hs * vs + ci - 1
                coef[mcu, bp] = zz[r, c]
    y = planes[0]
    cb = np.repeat(np.repeat(planes[1], vs, axis=0), hs, axis=1)
    cr = np.repeat(np.repeat(planes[2], vs, axis=0), hs, axis=1)
    y, cb, cr = (np.ascontiguousarray(p[:H, :W].astype(np.int16)) for p in (y, cb, cr))
    rgb = np.zeros((3, H, W), np.uint8)
    orc.gjdo_color_pixels(y.ctypes.data, cb.ctypes.data, cr.ctypes.data, H * W, rgb.ctypes.data)
    return coef.reshape(-1), rgb, (hs, vs)


def djpeg_nosmooth(data):
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(gen_synth.LIBJPEG, "lib"))
    out = subprocess.run([os.path.join(gen_synth.LIBJPEG, "bin", "djpeg"), "-nosmooth", "-pnm"], input=data, stdout=subprocess.PIPE,
                         check=True, env=env).stdout
    assert out[:2] == b"P6"
    parts = out.split(b"\n", 3)
    w, h = map(int, parts[1].split())
    return np.frombuffer(parts[3], np.uint8).reshape(h, w, 3).transpose(2, 0, 1)


def main():
    g = os.path.join(ROOT, "tests", "golden")
    os.makedirs(os.path.join(g, "sub"), exist_ok=True)
    orc = C.CDLL(os.path.join(ROOT, "oracle/_build/libgjd_oracle.so"))
    orc.gjdo_idct_block.argtypes = [C.c_void_p]
    orc.gjdo_color_pixels.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    zz_of_nat = np.ctypeslib.as_array((C.c_uint8 * 64).in_dll(orc, "gjdo_zigzag_of_natural")).astype(np.int64)
    files = {}
    for seed, w, h, q, samp, opt, rst in CASES:
        name = f"s{seed}_{w}x{h}_q{q}_{samp}_o{int(opt)}_rst{rst}"
        files[name] = encode(gen_synth.make_image(seed, w, h), q, samp, opt, rst)
    ref_img = "/root/reference/testing/image.jpeg"
    if os.path.exists(ref_img):
        files["ref_testing_image"] = open(ref_img, "rb").read()
    ref = {}
    for name, data in files.items():
        with open(os.path.join(g, "sub", name + ".jpg"), "wb") as f:
            f.write(data)
        coef, rgb, (hs, vs) = compose(data, orc, zz_of_nat)
        ref[name + "/coef"] = coef
        ref[name + "/rgb"] = rgb
        d = np.abs(rgb.astype(np.int32) - djpeg_nosmooth(data).astype(np.int32))
        print("%-44s %dx%d luma %dx%d  %6d bytes  vs djpeg -nosmooth: max |d| %d, mean %.3f" %
              (name, rgb.shape[2], rgb.shape[1], hs, vs, len(data), d.max(), d.mean()))
    np.savez_compressed(os.path.join(g, "ref_outputs_sub.npz"), **ref)


if __name__ == "__main__":
    main()
