import ctypes as C
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gpu-jpeg-decoder_b200")
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (PKG, os.path.join(ROOT, "tools"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _make(target_dir, *args):
    subprocess.run(["make", "-C", target_dir, *args], check=True, stdout=subprocess.DEVNULL)


class Oracle:
    """ctypes view of oracle/_build/libgjd_oracle.so (the CPU checker; test infrastructure only)."""

    def __init__(self):
        so = os.path.join(ROOT, "oracle", "_build", "libgjd_oracle.so")
        if not os.path.exists(so):
            _make(os.path.join(ROOT, "oracle"), "oracle")
        self.lib = C.CDLL(so)
        self.lib.gjdo_decode.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                         C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.gjdo_write_array.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p]
        self.lib.gjdo_write_array.restype = C.c_long
        self.lib.gjdo_color_plane.argtypes = [C.c_int, C.c_void_p]
        self.lib.gjdo_idct_block.argtypes = [C.c_void_p]
        self.lib.gjdo_dezigzag.argtypes = [C.c_void_p, C.c_void_p]

    def decode(self, data, want_coef=True, want_rgb=True, want_ycc=False):
        buf = C.cast(C.c_char_p(data), C.c_void_p)
        w, h = C.c_int(), C.c_int()
        rc = self.lib.gjdo_decode(buf, len(data), C.byref(w), C.byref(h), None, None, None)
        if rc != 0:
            return rc, None, None, None, 0, 0
        W, H = w.value, h.value
        nm = ((W + 7) // 8) * ((H + 7) // 8)
        coef = np.zeros(nm * 192, np.int16) if want_coef else None
        ycc = np.zeros(nm * 192, np.int16) if want_ycc else None
        rgb = np.zeros((3, H, W), np.uint8) if want_rgb else None
        rc = self.lib.gjdo_decode(buf, len(data), C.byref(w), C.byref(h),
                                  coef.ctypes.data if want_coef else None,
                                  ycc.ctypes.data if want_ycc else None,
                                  rgb.ctypes.data if want_rgb else None)
        return rc, coef, rgb, ycc, W, H


class HostSim:
    """tests/hostsim: the product's host/device functions stepped on the CPU (test scaffolding)."""

    def __init__(self):
        d = os.path.join(ROOT, "tests", "hostsim")
        bits = os.environ.get("GJD_LUT_BITS")       # A/B of the primary LUT width (csrc/gjd_types.h), default build otherwise
        so = os.path.join(d, "_build", "libgjd_hostsim%s.so" % ("_lut" + bits if bits else ""))
        srcs = [os.path.join(d, "hostsim.cpp"), os.path.join(PKG, "csrc", "gjd_parse.cpp")]
        deps = srcs + glob.glob(os.path.join(PKG, "csrc", "*.h")) + glob.glob(os.path.join(PKG, "csrc", "*.cuh"))
        if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(s) for s in deps):
            os.makedirs(os.path.dirname(so), exist_ok=True)
            subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                            "-Wno-unknown-pragmas", *(["-DGJD_LUT_BITS=" + bits] if bits else []),
                            "-o", so, *srcs], check=True)
        self.lib = C.CDLL(so)
        self.lib.hostsim_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                                            C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_void_p]
        self.lib.hostsim_color_plane.argtypes = [C.c_int, C.c_void_p]
        self.lib.hostsim_idct.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        self.lib.hostsim_idct_packed.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        self.lib.hostsim_color_plane_row8.argtypes = [C.c_int, C.c_void_p]
        self.lib.hostsim_pair_check.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_void_p]
        self.lib.hostsim_lut_check.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        self.lib.hostsim_parse.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        self.lib.hostsim_unstuff_device.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.hostsim_unstuff_device.restype = C.c_long
        self.lib.hostsim_unstuff_host.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.hostsim_unstuff_host.restype = C.c_long

    def parse(self, data):
        w, h = C.c_int32(), C.c_int32()
        rc = self.lib.hostsim_parse(C.cast(C.c_char_p(data), C.c_void_p), len(data), C.byref(w), C.byref(h))
        return rc, w.value, h.value

    def unstuff_both(self, data, misalign=0):
        """(host loop result, device-style result): each (length or negative status, bytes + restart table)."""
        buf = C.cast(C.c_char_p(data), C.c_void_p)
        a = np.zeros(len(data) + 128, np.uint8)
        b = np.zeros(len(data) + 128, np.uint8)
        ta = np.zeros(len(data) // 2 + 8, np.uint32)
        tb = np.zeros(len(data) // 2 + 8, np.uint32)
        na, nb = C.c_uint32(0), C.c_uint32(0)
        ra = self.lib.hostsim_unstuff_host(buf, len(data), a.ctypes.data, ta.ctypes.data, C.byref(na))
        rb = self.lib.hostsim_unstuff_device(buf, len(data), misalign, b.ctypes.data, tb.ctypes.data, C.byref(nb))
        return ((ra, a[: max(ra, 0)].tobytes() + ta[: na.value].tobytes() if ra >= 0 else b""),
                (rb, b[: max(rb, 0)].tobytes() + tb[: nb.value].tobytes() if rb >= 0 else b""))

    def decode(self, data, sub_log2=0, ncoef=None):
        """ncoef: int16 coefficients to return (default: the 4:4:4 count, ceil(W/8)*ceil(H/8)*192; chroma-subsampled
        files have 64 * blocks-per-MCU * MCUs, in scan order)."""
        rc, W, H = self.parse(data)
        if rc != 0:
            return rc, None, None, None
        # room for the subsampled layouts too (up to 6 blocks per 16x16 MCU < 3 blocks per 8x8 one)
        nm = ((W + 15) // 8) * ((H + 15) // 8)
        coef = np.zeros(nm * 192, np.int16)      # grayscale files fill the first nm * 64
        rgb = np.zeros((3, H, W), np.uint8)
        stats = np.zeros(9, np.uint32)   # [8] = repeat decodes cut short at a checkpoint
        w, h = C.c_int32(), C.c_int32()
        rc = self.lib.hostsim_decode(C.cast(C.c_char_p(data), C.c_void_p), len(data), sub_log2, coef.ctypes.data,
                                     rgb.ctypes.data, C.byref(w), C.byref(h), stats.ctypes.data)
        return rc, coef[: (ncoef if ncoef is not None else ((W + 7) // 8) * ((H + 7) // 8) * 192)], rgb, stats


@pytest.fixture(scope="session")
def oracle():
    return Oracle()


@pytest.fixture(scope="session")
def hostsim():
    return HostSim()


def golden_files():
    out = {}
    for f in sorted(glob.glob(os.path.join(GOLDEN, "images", "*.jpg")) + glob.glob(os.path.join(GOLDEN, "synth", "*.jpg"))):
        name = os.path.basename(os.path.dirname(f)) + "/" + os.path.basename(f)[:-4]
        with open(f, "rb") as fh:
            out[name] = fh.read()
    return out


@pytest.fixture(scope="session")
def golden():
    return golden_files()


@pytest.fixture(scope="session")
def rst_golden():
    out = {}
    for f in sorted(glob.glob(os.path.join(GOLDEN, "rst", "*.jpg"))):
        with open(f, "rb") as fh:
            out[os.path.basename(f)[:-4]] = fh.read()
    return out


@pytest.fixture(scope="session")
def gray_golden():
    out = {}
    for f in sorted(glob.glob(os.path.join(GOLDEN, "gray", "*.jpg"))):
        with open(f, "rb") as fh:
            out[os.path.basename(f)[:-4]] = fh.read()
    return out


@pytest.fixture(scope="session")
def sub_golden():
    out = {}
    for f in sorted(glob.glob(os.path.join(GOLDEN, "sub", "*.jpg"))):
        with open(f, "rb") as fh:
            out[os.path.basename(f)[:-4]] = fh.read()
    return out


@pytest.fixture(scope="session")
def ref_outputs_sub():
    return np.load(os.path.join(GOLDEN, "ref_outputs_sub.npz"))


@pytest.fixture(scope="session")
def ref_outputs_gray():
    return np.load(os.path.join(GOLDEN, "ref_outputs_gray.npz"))


@pytest.fixture(scope="session")
def ref_outputs_rst():
    return np.load(os.path.join(GOLDEN, "ref_outputs_rst.npz"))


@pytest.fixture(scope="session")
def ref_outputs():
    return np.load(os.path.join(GOLDEN, "ref_outputs.npz"))


@pytest.fixture(scope="session")
def ground_truth():
    return np.load(os.path.join(GOLDEN, "ground_truth.npz"))


@pytest.fixture(scope="session")
def decoder():
    """A gjd_ctx on cuda:0 through the C ABI.  Fails (does not skip) if libgjd.so is missing."""
    import gjd
    d = gjd.Decoder(0)
    yield d
    d.close()
