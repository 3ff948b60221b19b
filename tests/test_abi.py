"""The C-ABI library loads on a machine without a GPU, exports every symbol include/gjd.h declares, and
refuses to create a context instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import numpy as np

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "gjd.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gjd_[a-z_0-9]+)\s*\(", src)))


def test_exports_every_declared_symbol():
    import gjd
    L = gjd.lib()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"libgjd.so does not export {n}"


def test_version_and_strerror():
    import gjd
    L = gjd.lib()
    assert b"sm_100a" in L.gjd_version()
    assert L.gjd_strerror(-7).startswith(b"no usable CUDA device")


def test_peek_is_host_only(golden):
    import gjd
    rc, w, h = gjd.peek(golden["images/8_401x363"])
    assert (rc, w, h) == (0, 401, 363)
    rc, w, h = gjd.peek(b"\xff\xd8\xff\xd9")
    assert rc == gjd.E_FORMAT


def test_no_cpu_fallback_without_device():
    import gjd
    import torch
    if torch.cuda.is_available():
        return
    ctx = C.c_void_p()
    rc = gjd.lib().gjd_create(C.byref(ctx), 0, 0)
    assert rc == gjd.E_NODEVICE and not ctx.value


def test_array_writer_format(golden, ground_truth, tmp_path):
    import hashlib, json, gjd
    from conftest import GOLDEN
    sha = json.load(open(os.path.join(GOLDEN, "array_sha256.json")))
    for stem, want in sha.items():
        p = str(tmp_path / (stem + ".array"))
        gjd.write_array(p, ground_truth[stem])
        assert hashlib.sha256(open(p, "rb").read()).hexdigest() == want


def test_throughput_driver_without_gpu_fails_loudly():
    """No CPU fallback: on a machine without a CUDA device the benchmark driver reports it and exits non-zero."""
    import subprocess
    exe = os.path.join(ROOT, "gjd-decoder", "gjd_bench")
    if not os.path.exists(exe):
        import pytest
        pytest.skip("gjd_bench not built")
    assert subprocess.run([exe], stderr=subprocess.PIPE).returncode == 1
    import gjd
    ctx = C.c_void_p()
    if gjd.lib().gjd_create(C.byref(ctx), 0, 0) == 0:      # a GPU is present: nothing to check here
        gjd.lib().gjd_destroy(ctx)
        return
    p = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "images")], stderr=subprocess.PIPE, text=True)
    assert p.returncode == 3 and "CUDA device" in p.stderr


def test_item_struct_matches_header():
    """The ctypes mirror of gjd_item has the header's fields, in order, and its size (64 bytes)."""
    import gjd
    src = open(os.path.join(ROOT, "include", "gjd.h")).read()
    body = src[src.index("typedef struct gjd_item {"):src.index("} gjd_item;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = re.findall(r"\b(\w+);", body)
    assert fields == [f[0] for f in gjd.Item._fields_]
    assert C.sizeof(gjd.Item) == 64


def test_ppm_writer_roundtrip(ground_truth, tmp_path):
    """gjd_write_ppm -> the reader logic of the reference's jpeglib-implementation/process_ppm.py:4-41 (P6 header,
    optional comment lines, dimensions, maxval, raw interleaved bytes split into R, G, B) gives the planes back."""
    import gjd
    rng = np.random.default_rng(5)
    cases = [ground_truth[k] for k in ground_truth.files[:2]] + [rng.integers(0, 256, (3, 7, 13), dtype=np.uint8),
                                                                  rng.integers(0, 256, (3, 1, 1), dtype=np.uint8)]
    for i, rgb in enumerate(cases):
        p = str(tmp_path / ("x%d.ppm" % i))
        gjd.write_ppm(p, rgb)
        with open(p, "rb") as f:
            assert f.readline().decode().strip() == "P6"
            while True:
                line = f.readline().decode().strip()
                if not line.startswith("#"):
                    break
            width, height = map(int, line.split())
            assert int(f.readline().decode().strip()) == 255
            raw = list(f.read())
        assert (height, width) == rgb.shape[1:] and len(raw) == 3 * width * height
        for c in range(3):
            assert raw[c::3] == rgb[c].ravel().tolist()
    assert gjd.lib().gjd_write_ppm(str(tmp_path / "nodir" / "x.ppm").encode(), 1, 1, cases[-1].ctypes.data) == gjd.E_IO
    assert gjd.lib().gjd_write_ppm(None, 1, 1, cases[-1].ctypes.data) == gjd.E_ARG
