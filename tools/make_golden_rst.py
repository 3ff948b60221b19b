"""Generate tests/golden/rst/*.jpg + tests/golden/ref_outputs_rst.npz (run in the build container).

Restart markers (DRI/RSTn) are outside the reference's parser (cpp-decoder/src/parser.cpp:84-96 keeps
FF Dn as scan data), so the pin is self-consistency: `cjpeg -restart N` emits the same
quantised coefficients as the same call without it, hence a decoder with the reference's arithmetic must
give, for the restart file, exactly what the UNMODIFIED reference (oracle/_ref/libgjd_ref.so) gives for
the twin encoded without restarts.  The npz holds those reference outputs ("<name>/coef", "<name>/rgb");
the twin files themselves are not kept.
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

CASES = [  # (seed, w, h, quality, optimize, cjpeg -restart argument)
    (21, 97, 131, 90, False, "1"),      # one MCU row = 13 MCUs: intervals cross tiles at odd offsets
    (22, 256, 256, 90, True, "1B"),     # every MCU is its own interval
    (23, 401, 363, 75, False, "7B"),    # 7 MCUs: never aligned with the 32-MCU tiles
    (24, 640, 480, 90, False, "2"),     # 160 MCUs = 5 tiles per interval
    (25, 33, 17, 95, False, "100B"),    # interval longer than the image: DRI present, no RSTn in the scan
    (26, 520, 200, 85, True, "40B"),    # 40 MCUs: base taken from inside another tile
    (27, 1024, 64, 90, False, "9B"),    # more than 8 intervals per tile row: RSTn numbering wraps
]


def main():
    g = os.path.join(ROOT, "tests", "golden")
    os.makedirs(os.path.join(g, "rst"), exist_ok=True)
    lib = C.CDLL(os.path.join(ROOT, "oracle/_ref/libgjd_ref.so"))
    lib.gjdref_decode.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                  C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_long), C.POINTER(C.POINTER(C.c_int32))]
    lib.gjdref_quant.argtypes = [C.c_char_p, C.c_void_p]
    lib.gjdref_free.argtypes = [C.c_void_p]
    ref = {}
    for seed, w, h, q, opt, rst in CASES:
        img = gen_synth.make_image(seed, w, h)
        name = f"r{seed}_{w}x{h}_q{q}_o{int(opt)}_rst{rst}"
        data = gen_synth.encode(img, q, opt, restart=rst)
        twin = gen_synth.encode(img, q, opt)
        assert data != twin and (b"\xff\xdd" in data)
        with open(os.path.join(g, "rst", name + ".jpg"), "wb") as f:
            f.write(data)
        with tempfile.NamedTemporaryFile(suffix=".jpg") as tf:
            tf.write(twin)
            tf.flush()
            coef, px = ref_decode(lib, tf.name)
        ref[name + "/coef"] = coef
        ref[name + "/rgb"] = px
        print(name, len(data), "bytes,", data.count(b"\xff\xd0") , "x RST0")
    np.savez_compressed(os.path.join(g, "ref_outputs_rst.npz"), **ref)


if __name__ == "__main__":
    main()
