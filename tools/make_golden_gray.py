"""Generate tests/golden/gray/*.jpg + tests/golden/ref_outputs_gray.npz (run in the build container).

Single-component files are outside the reference's parser (it wants three components), so the pin is built
from the reference's own pieces:
  coefficients  `cjpeg -grayscale` codes the same Y plane with the same luminance table as the colour
                encoding of the same picture, so the file's blocks must equal the Y blocks the UNMODIFIED
                reference (oracle/_ref/libgjd_ref.so tap) reports for the colour twin;
  pixels        those blocks through the oracle's dezigzag + IDCT + colour formula with Cb = Cr = 0
                (oracle/gjd_oracle.c, itself pinned to the compiled reference by tests/test_oracle.py).
"<name>/coef" int16 [block*64 + k] (zigzag, absolute DC), "<name>/rgb" (3,H,W) uint8.
"""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth  # noqa: E402
from make_golden import ref_decode  # noqa: E402

CASES = [  # (seed, w, h, quality, optimize, cjpeg -restart argument or 0)
    (31, 64, 64, 90, False, 0),
    (32, 401, 363, 75, True, 0),       # ragged edges, optimised tables
    (33, 1000, 24, 90, False, 0),      # more than one 96-block tile per block row
    (34, 97, 131, 95, False, "1"),     # restart every block row (13 blocks)
    (35, 520, 200, 85, False, "100B"), # restart interval of 100 blocks: spans tiles
    (36, 1, 1, 90, False, 0),
]


def main():
    g = os.path.join(ROOT, "tests", "golden")
    os.makedirs(os.path.join(g, "gray"), exist_ok=True)
    lib = C.CDLL(os.path.join(ROOT, "oracle/_ref/libgjd_ref.so"))
    lib.gjdref_decode.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                  C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_long), C.POINTER(C.POINTER(C.c_int32))]
    lib.gjdref_quant.argtypes = [C.c_char_p, C.c_void_p]
    lib.gjdref_free.argtypes = [C.c_void_p]
    orc = C.CDLL(os.path.join(ROOT, "oracle/_build/libgjd_oracle.so"))
    orc.gjdo_dezigzag.argtypes = [C.c_void_p, C.c_void_p]
    orc.gjdo_idct_block.argtypes = [C.c_void_p]
    orc.gjdo_color.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lut = np.zeros((512, 3), np.uint8)          # colour formula at Cb = Cr = 0 for every sample value
    for y in range(-256, 256):
        r, gg, b = C.c_int(), C.c_int(), C.c_int()
        orc.gjdo_color(y, 0, 0, C.byref(r), C.byref(gg), C.byref(b))
        lut[y + 256] = (r.value, gg.value, b.value)
    ref = {}
    for seed, w, h, q, opt, rst in CASES:
        img = gen_synth.make_image(seed, w, h)
        name = f"g{seed}_{w}x{h}_q{q}_o{int(opt)}_rst{rst}"
        data = gen_synth.encode(img, q, opt, restart=rst, grayscale=True)
        twin = gen_synth.encode(img, q, opt)
        with open(os.path.join(g, "gray", name + ".jpg"), "wb") as f:
            f.write(data)
        with tempfile.NamedTemporaryFile(suffix=".jpg") as tf:
            tf.write(twin)
            tf.flush()
            coef, _ = ref_decode(lib, tf.name)
            qt = (C.c_uint8 * 128)()
            assert lib.gjdref_quant(tf.name.encode(), qt) == 0
        qy = np.frombuffer(qt, np.uint8)[:64].astype(np.int32)
        ycoef = coef.reshape(-1, 3, 64)[:, 0, :].astype(np.int32)          # the Y blocks of the twin
        bw, bh = (w + 7) // 8, (h + 7) // 8
        planes = np.zeros((bh * 8, bw * 8), np.int32)
        for i in range(bw * bh):
            zz = np.ascontiguousarray(ycoef[i] * qy, np.int32)
            nat = np.zeros(64, np.int32)
            orc.gjdo_dezigzag(zz.ctypes.data, nat.ctypes.data)
            orc.gjdo_idct_block(nat.ctypes.data)
            by, bx = divmod(i, bw)
            planes[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8] = nat.reshape(8, 8)
        px = lut[planes[:h, :w] + 256]                                     # (h, w, 3)
        ref[name + "/coef"] = ycoef.astype(np.int16).reshape(-1)
        ref[name + "/rgb"] = np.ascontiguousarray(px.transpose(2, 0, 1))
        print(name, len(data), "bytes", "R!=G pixels:", int(np.count_nonzero(px[..., 0] != px[..., 1])))
    np.savez_compressed(os.path.join(g, "ref_outputs_gray.npz"), **ref)


if __name__ == "__main__":
    main()
