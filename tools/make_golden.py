"""Generate tests/golden/* from the reference (run in the build container, where /root/reference exists).

  tests/golden/images/*.jpg        the reference's testing/images fixtures (data, copied verbatim)
  tests/golden/synth/*.jpg         small synthetic JPEGs (tools/gen_synth.py + vendored cjpeg): edge sizes,
                                   optimised tables, several qualities
  tests/golden/ground_truth.npz    <stem> -> (3,H,W) uint8 parsed from the reference's committed
                                   testing/ground_truth/<stem>.array; 4_800x600 (absent from the
                                   mirror) is regenerated with the unmodified reference CLI
                                   (oracle/_ref/decoder) and stored under "regen/4_800x600"
  tests/golden/array_sha256.json   sha256 of the reference's committed testing/ground_truth/*.array files
  tests/golden/ref_outputs.npz     for every image above, outputs of the UNMODIFIED reference compiled
                                   into oracle/_ref/libgjd_ref.so: "<name>/coef" quantised zigzag
                                   coefficients int16 [mcu*192+comp*64+k] (tapped dequantised values
                                   divided by the quant table), "<name>/rgb" (3,H,W) uint8

Usage: python tools/make_golden.py
"""
import ctypes as C
import glob
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_synth  # noqa: E402

SYNTH = [  # (seed, w, h, quality, optimize)
    (1, 8, 8, 90, False), (2, 16, 8, 75, False), (3, 33, 17, 90, False), (4, 64, 64, 50, False),
    (5, 97, 131, 95, True), (6, 256, 256, 90, True), (7, 401, 363, 100, False), (8, 512, 384, 25, False),
    (9, 640, 480, 90, False), (10, 1, 1, 90, False), (11, 7, 300, 85, True), (12, 300, 7, 85, False),
]


def parse_array(path):
    with open(path) as f:
        h, w = map(int, f.readline().split())
        planes = [np.array(f.readline().split(), dtype=np.int64) for _ in range(3)]
    return np.stack([p.reshape(h, w) for p in planes]).astype(np.uint8)


def ref_decode(lib, path):
    w, h = C.c_int(), C.c_int()
    deq, rgb, n = C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)(), C.c_long()
    rc = lib.gjdref_decode(path.encode(), C.byref(w), C.byref(h), C.byref(deq), C.byref(n), C.byref(rgb))
    assert rc == 0
    W, H = w.value, h.value
    d = np.ctypeslib.as_array(deq, (n.value,)).copy().reshape(-1, 3, 64)
    px = np.ctypeslib.as_array(rgb, (3 * W * H,)).copy().reshape(3, H, W)
    lib.gjdref_free(deq)
    lib.gjdref_free(rgb)
    q = (C.c_uint8 * 128)()
    assert lib.gjdref_quant(path.encode(), q) == 0
    q = np.frombuffer(q, np.uint8).astype(np.int32).reshape(2, 64)
    qq = np.stack([q[0], q[1], q[1]])           # Y -> table 0, Cb/Cr -> table 1 (parser.cpp:163-165)
    assert np.all(d % qq[None] == 0)
    coef = (d // qq[None]).astype(np.int16).reshape(-1)
    assert px.min() >= 0 and px.max() <= 255
    return coef, px.astype(np.uint8)


def main():
    g = os.path.join(ROOT, "tests", "golden")
    os.makedirs(os.path.join(g, "images"), exist_ok=True)
    os.makedirs(os.path.join(g, "synth"), exist_ok=True)
    for f in sorted(glob.glob(os.path.join(REF, "testing/images/*.jpg"))):
        shutil.copyfile(f, os.path.join(g, "images", os.path.basename(f)))
    for seed, w, h, q, opt in SYNTH:
        data = gen_synth.encode(gen_synth.make_image(seed, max(w, 1), max(h, 1)), q, opt)
        with open(os.path.join(g, "synth", f"s{seed}_{w}x{h}_q{q}_o{int(opt)}.jpg"), "wb") as f:
            f.write(data)

    truth = {}
    for f in sorted(glob.glob(os.path.join(REF, "testing/ground_truth/*.array"))):
        stem = os.path.basename(f)[:-6]
        if os.path.exists(os.path.join(g, "images", stem + ".jpg")):
            truth[stem] = parse_array(f)
    # regenerate the missing 4_800x600 truth with the unmodified reference CLI
    with tempfile.TemporaryDirectory() as td:
        os.makedirs(os.path.join(td, "testing", "cpp_output_arrays"))
        os.makedirs(os.path.join(td, "cpp-decoder"))
        subprocess.run([os.path.join(ROOT, "oracle/_ref/decoder"), os.path.join(g, "images", "4_800x600.jpg")],
                       cwd=os.path.join(td, "cpp-decoder"), check=True, stdout=subprocess.DEVNULL)
        truth["regen/4_800x600"] = parse_array(os.path.join(td, "testing/cpp_output_arrays/4_800x600.array"))
    np.savez_compressed(os.path.join(g, "ground_truth.npz"), **truth)
    # byte-level pin of the `.array` text format: sha256 of the reference's committed files
    import hashlib, json
    sha = {}
    for f in sorted(glob.glob(os.path.join(REF, "testing/ground_truth/*.array"))):
        stem = os.path.basename(f)[:-6]
        if stem in truth:
            sha[stem] = hashlib.sha256(open(f, "rb").read()).hexdigest()
    with open(os.path.join(g, "array_sha256.json"), "w") as f:
        json.dump(sha, f, indent=1, sort_keys=True)

    lib = C.CDLL(os.path.join(ROOT, "oracle/_ref/libgjd_ref.so"))
    lib.gjdref_decode.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                  C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_long), C.POINTER(C.POINTER(C.c_int32))]
    lib.gjdref_quant.argtypes = [C.c_char_p, C.c_void_p]
    lib.gjdref_free.argtypes = [C.c_void_p]
    ref = {}
    for f in sorted(glob.glob(os.path.join(g, "images", "*.jpg")) + glob.glob(os.path.join(g, "synth", "*.jpg"))):
        name = os.path.basename(os.path.dirname(f)) + "/" + os.path.basename(f)[:-4]
        coef, px = ref_decode(lib, f)
        ref[name + "/coef"] = coef
        ref[name + "/rgb"] = px
        print(name, px.shape, "nonzero coefs", int(np.count_nonzero(coef)))
    np.savez_compressed(os.path.join(g, "ref_outputs.npz"), **ref)
    # the reference's only coefficient-level known answer: legacy_versions/cudaU-implementation/gold.txt
    gold = [[-268, -68, -8], [-19, 0, 3], [-13, 0, -2], [-7, 0, 1], [0, 0, 2], [-1, 0, 0], [1, 0, 0], [7, 0, 0], [0, 0, 0], [-7, 0, 0]]
    for name in list(ref):
        if name.endswith("/coef"):
            c = ref[name].reshape(-1, 3, 64)[0]
            if [[int(c[0][k]), int(c[1][k]), int(c[2][k])] for k in range(10)] == gold:
                print("gold.txt matches the first MCU of", name)


if __name__ == "__main__":
    main()
