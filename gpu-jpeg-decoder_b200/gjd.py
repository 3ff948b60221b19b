"""ctypes shim over libgjd.so (include/gjd.h).

This is the Python mirror of the reference-facing interface: the reference's callers are C++
(`cuda-decoder/main.cu:32-39`, `cuda-decoder/benchmark_thoughput/benchmark.cu:49-93`) and its
harness is Python driving the CLI (`testing/compare.py:37-64`); the tests and bench.py use this shim
to call the same C ABI those callers would bind.  There is no Python or CPU decode path here: if the
shared library is missing the import of `lib()` raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GJD_LIB") or os.path.join(_HERE, "build", "libgjd.so")   # GJD_LIB: A/B builds

PLANAR, INTERLEAVED, INTERLEAVED_I16 = 0, 1, 2
PARTITION_BYTE_WEIGHT = 5          # GJD_PARTITION_BYTE_WEIGHT (include/gjd.h)
OK, E_ARG, E_CUDA, E_NOMEM, E_FORMAT, E_UNSUPPORTED, E_CORRUPT, E_NODEVICE, E_IO = 0, -1, -2, -3, -4, -5, -6, -7, -8


class Item(C.Structure):
    _fields_ = [
        ("jpg", C.c_void_p),
        ("len", C.c_size_t),
        ("rgb", C.c_void_p),
        ("plane_stride", C.c_size_t),
        ("row_pitch", C.c_size_t),
        ("rgb_on_device", C.c_int32),
        ("width", C.c_int32),
        ("height", C.c_int32),
        ("status", C.c_int32),
        ("layout", C.c_int32),
        ("reserved", C.c_int32),
    ]


_lib = None


def lib():
    """Load libgjd.so; fail loudly when the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()). "
            "gjd has no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int32
    L.gjd_create.argtypes = [C.POINTER(vp), C.c_int, sz]
    L.gjd_destroy.argtypes = [vp]
    L.gjd_destroy.restype = None
    L.gjd_peek.argtypes = [vp, sz, C.POINTER(i32), C.POINTER(i32)]
    L.gjd_decode_batch.argtypes = [vp, C.POINTER(Item), i32]
    L.gjd_batch_prepare.argtypes = [vp, C.POINTER(Item), i32, C.POINTER(vp)]
    L.gjd_batch_run.argtypes = [vp, vp]
    L.gjd_batch_sync.argtypes = [vp, vp]
    L.gjd_batch_status.argtypes = [vp, vp, i32]
    L.gjd_batch_fetch.argtypes = [vp, vp, i32, vp]
    L.gjd_batch_free.argtypes = [vp, vp]
    L.gjd_batch_free.restype = None
    L.gjd_batch_info.argtypes = [vp, vp, C.POINTER(C.c_uint64)]
    L.gjd_decode_coeffs.argtypes = [vp, vp, sz, vp]
    L.gjd_stage_times.argtypes = [vp, C.POINTER(C.c_float)]
    L.gjd_test_color.argtypes = [vp, vp, vp, vp, sz, vp, vp, vp]
    L.gjd_test_idct.argtypes = [vp, vp, sz, vp, vp]
    L.gjd_idct_blocks_static.argtypes = [C.c_int, vp, sz, vp, vp, C.c_char_p, sz]
    L.gjd_host_alloc.argtypes = [sz]
    L.gjd_host_alloc.restype = vp
    L.gjd_host_free.argtypes = [vp]
    L.gjd_host_free.restype = None
    L.gjd_write_array.argtypes = [C.c_char_p, i32, i32, vp]
    L.gjd_write_ppm.argtypes = [C.c_char_p, i32, i32, vp]
    L.gjd_partition.argtypes = [vp, i32, i32, vp, vp]
    L.gjd_partition_weights.argtypes = [C.POINTER(Item), i32, vp]
    L.gjd_partition_costs.argtypes = [C.POINTER(Item), i32, C.c_uint32, vp]
    L.gjd_strerror.argtypes = [C.c_int]
    L.gjd_strerror.restype = C.c_char_p
    L.gjd_last_cuda_error.argtypes = [vp]
    L.gjd_last_cuda_error.restype = C.c_char_p
    L.gjd_version.restype = C.c_char_p
    _lib = L
    return L


class GjdError(RuntimeError):
    def __init__(self, status, detail=""):
        self.status = status
        msg = lib().gjd_strerror(status).decode()
        super().__init__(f"gjd status {status}: {msg} {detail}".strip())


def sof_sampling(data: bytes):
    """(components, luma h, luma v) from the SOF0 segment (host-side helper for sizing coefficient buffers)."""
    i = 2
    while i + 4 <= len(data):
        if data[i] != 0xFF:
            break
        m = data[i + 1]
        if m == 0xFF:
            i += 1
            continue
        seg = (data[i + 2] << 8) | data[i + 3]
        if m == 0xC0:
            nf = data[i + 9]
            return nf, (data[i + 11] >> 4) if nf == 3 else 1, (data[i + 11] & 15) if nf == 3 else 1
        if m == 0xDA:
            break
        i += 2 + seg
    return 3, 1, 1


def peek(data: bytes):
    w, h = C.c_int32(), C.c_int32()
    rc = lib().gjd_peek(C.cast(C.c_char_p(data), C.c_void_p), len(data), C.byref(w), C.byref(h))
    return rc, w.value, h.value


class PinnedBuffer:
    """Page-locked host memory from gjd_host_alloc, exposed as a uint8 numpy array."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        self.ptr = lib().gjd_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise MemoryError("gjd_host_alloc failed")
        self.array = np.ctypeslib.as_array((C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr))[: self.nbytes]

    def free(self):
        if self.ptr:
            self.array = None
            lib().gjd_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Decoder:
    """One gjd_ctx (one host thread, one CUDA device)."""

    def __init__(self, device=0, scratch_bytes=0):
        self._ctx = C.c_void_p()
        rc = lib().gjd_create(C.byref(self._ctx), device, scratch_bytes)
        if rc != OK:
            raise GjdError(rc)

    def close(self):
        if self._ctx:
            lib().gjd_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _detail(self):
        return lib().gjd_last_cuda_error(self._ctx).decode()

    # -- end-to-end ---------------------------------------------------------------------------
    def make_items(self, files):
        """Item array for a list of `bytes` (kept alive by the returned holder list)."""
        n = len(files)
        items = (Item * n)()
        hold = []
        for i, d in enumerate(files):
            buf = C.c_char_p(d)
            hold.append(buf)
            items[i].jpg = C.cast(buf, C.c_void_p)
            items[i].len = len(d)
        return items, hold

    def make_items_pinned(self, files):
        """The same, with the JPEG bytes packed back to back in ONE page-locked buffer (gjd_host_alloc): the library
        then DMAs straight from it.  Returns (items, PinnedBuffer)."""
        n = len(files)
        items = (Item * max(n, 1))()
        pin = PinnedBuffer(sum(len(d) for d in files) + 16)
        off = 0
        for i, d in enumerate(files):
            pin.array[off:off + len(d)] = np.frombuffer(d, np.uint8)
            items[i].jpg = pin.ptr + off
            items[i].len = len(d)
            off += len(d)
        return items, pin

    def decode(self, files, layout=PLANAR, pinned=False):
        """Decode a list of JPEG `bytes`; returns (list of arrays or None, list of status).
        Arrays are uint8 (3,H,W) for the planar layout, uint8 (H,W,3) for the interleaved one and int16 (H,W,3)
        for INTERLEAVED_I16 (element type and order of the reference's outputChannels)."""
        n = len(files)
        items, hold = self.make_items_pinned(files) if pinned else self.make_items(files)
        outs = []
        for i, d in enumerate(files):
            rc, w, h = peek(d)
            items[i].layout = layout
            if rc == OK:
                a = np.empty((3, h, w) if layout == PLANAR else (h, w, 3), np.int16 if layout == INTERLEAVED_I16 else np.uint8)
                items[i].rgb = a.ctypes.data
                outs.append(a)
            else:
                outs.append(None)
        rc = lib().gjd_decode_batch(self._ctx, items, n)
        if rc != OK:
            raise GjdError(rc, self._detail())
        st = [items[i].status for i in range(n)]
        return [o if s == OK else None for o, s in zip(outs, st)], st

    def decode_batch_raw(self, items, n):
        rc = lib().gjd_decode_batch(self._ctx, items, n)
        if rc != OK:
            raise GjdError(rc, self._detail())

    def decode_coeffs(self, data: bytes):
        """Quantised coefficients (gjd_decode_coeffs): 4:4:4 files mcu*192 + comp*64 + k; single-component files
        block*64 + k; chroma-subsampled files line*64 + k with line = mcu*(hs*vs + 2) + block-in-MCU (scan order)."""
        rc, w, h = peek(data)
        if rc != OK:
            raise GjdError(rc)
        nc, hs, vs = sof_sampling(data)
        nm = ((w + 8 * hs - 1) // (8 * hs)) * ((h + 8 * vs - 1) // (8 * vs))
        out = np.empty(nm * (hs * vs + 2) * 64 if nc == 3 else nm * 192, np.int16)
        rc = lib().gjd_decode_coeffs(self._ctx, C.cast(C.c_char_p(data), C.c_void_p), len(data), out.ctypes.data)
        if rc != OK:
            raise GjdError(rc, self._detail())
        return out, w, h

    def stage_times(self):
        ms = (C.c_float * 4)()
        rc = lib().gjd_stage_times(self._ctx, ms)
        if rc != OK:
            raise GjdError(rc, self._detail())
        return list(ms)

    # -- resident batch -----------------------------------------------------------------------
    def prepare(self, files, layout=PLANAR):
        items, hold = self.make_items(files)
        for i in range(len(files)):
            items[i].layout = layout
        b = C.c_void_p()
        rc = lib().gjd_batch_prepare(self._ctx, items, len(files), C.byref(b))
        if rc != OK:
            raise GjdError(rc, self._detail())
        return ResidentBatch(self, b, items, len(files))

    # -- test hooks ---------------------------------------------------------------------------
    def test_color(self, y, cb, cr):
        y = np.ascontiguousarray(y, np.int16)
        cb = np.ascontiguousarray(cb, np.int16)
        cr = np.ascontiguousarray(cr, np.int16)
        n = y.size
        r, g, b = (np.empty(n, np.uint8) for _ in range(3))
        rc = lib().gjd_test_color(self._ctx, y.ctypes.data, cb.ctypes.data, cr.ctypes.data, n,
                                  r.ctypes.data, g.ctypes.data, b.ctypes.data)
        if rc != OK:
            raise GjdError(rc, self._detail())
        return r, g, b

    def test_idct(self, coef_zz, quant_zz):
        coef_zz = np.ascontiguousarray(coef_zz, np.int16).reshape(-1, 64)
        q = np.ascontiguousarray(quant_zz, np.uint8)
        out = np.empty_like(coef_zz)
        rc = lib().gjd_test_idct(self._ctx, coef_zz.ctypes.data, coef_zz.shape[0], q.ctypes.data, out.ctypes.data)
        if rc != OK:
            raise GjdError(rc, self._detail())
        return out


class ResidentBatch:
    def __init__(self, dec, handle, items, n):
        self.dec, self.h, self.items, self.n = dec, handle, items, n

    def run(self):
        rc = lib().gjd_batch_run(self.dec._ctx, self.h)
        if rc != OK:
            raise GjdError(rc, self.dec._detail())

    def sync(self):
        return lib().gjd_batch_sync(self.dec._ctx, self.h)

    def status(self, i):
        return lib().gjd_batch_status(self.dec._ctx, self.h, i)

    def info(self):
        v = (C.c_uint64 * 5)()
        lib().gjd_batch_info(self.dec._ctx, self.h, v)
        return dict(pixels=v[0], padded_pixels=v[1], scan_bytes=v[2], subseqs=v[3], launches=v[4])

    def fetch(self, i):
        w, h = self.items[i].width, self.items[i].height
        layout = self.items[i].layout
        out = np.empty((3, h, w) if layout == PLANAR else (h, w, 3), np.int16 if layout == INTERLEAVED_I16 else np.uint8)
        rc = lib().gjd_batch_fetch(self.dec._ctx, self.h, i, out.ctypes.data)
        if rc != OK:
            raise GjdError(rc, self.dec._detail())
        return out

    def free(self):
        if self.h:
            lib().gjd_batch_free(self.dec._ctx, self.h)
            self.h = None


STATIC_IDCT_MAX_BLOCKS = 8192   # GJD_STATIC_IDCT_MAX_BLOCKS (include/gjd.h)


def idct_blocks_static(coef_zz, quant_zz, device=0, out=None):
    """gjd_idct_blocks_static: Decoder.test_idct without a context and without any device allocation (the scratch is a
    static device array inside libgjd.so).  coef_zz / out may also be integer device addresses (then pass nblocks rows
    through `coef_zz=(ptr, nblocks)` and `out=ptr`); numpy arrays are host memory."""
    q = np.ascontiguousarray(quant_zz, np.uint8)
    if isinstance(coef_zz, tuple):
        src, n = coef_zz
        result = dst = out
    else:
        coef_zz = np.ascontiguousarray(coef_zz, np.int16).reshape(-1, 64)
        src, n = coef_zz.ctypes.data, coef_zz.shape[0]
        result = np.empty_like(coef_zz) if out is None else out
        dst = result if isinstance(result, int) else result.ctypes.data
    err = C.create_string_buffer(256)
    rc = lib().gjd_idct_blocks_static(device, src, n, q.ctypes.data, dst, err, len(err))
    if rc != OK:
        raise GjdError(rc, err.value.decode())
    return result


def write_array(path, rgb):
    """rgb: (3,H,W) uint8."""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    rc = lib().gjd_write_array(path.encode(), rgb.shape[2], rgb.shape[1], rgb.ctypes.data)
    if rc != OK:
        raise GjdError(rc)


def write_ppm(path, rgb):
    """rgb: (3,H,W) uint8 -> binary P6."""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    rc = lib().gjd_write_ppm(path.encode(), rgb.shape[2], rgb.shape[1], rgb.ctypes.data)
    if rc != OK:
        raise GjdError(rc)


def partition(files, parts, byte_weight=0):
    """Greedy-LPT assignment of JPEG `bytes` objects to `parts` devices (gjd_partition over
    gjd_partition_weights, or over gjd_partition_costs when byte_weight > 0: padded pixels +
    byte_weight * compressed length).  Returns (part_of int32[n], load uint64[parts], weight uint64[n]);
    load and weight are in the unit the partition balanced."""
    n = len(files)
    items = (Item * max(n, 1))()
    hold = []
    for i, d in enumerate(files):
        buf = C.c_char_p(d)
        hold.append(buf)
        items[i].jpg = C.cast(buf, C.c_void_p)
        items[i].len = len(d)
    weight = np.zeros(max(n, 1), np.uint64)
    part_of = np.full(max(n, 1), -1, np.int32)
    load = np.zeros(parts, np.uint64)
    if byte_weight:
        rc = lib().gjd_partition_costs(items, n, int(byte_weight), weight.ctypes.data)
    else:
        rc = lib().gjd_partition_weights(items, n, weight.ctypes.data)
    if rc == OK:
        rc = lib().gjd_partition(weight.ctypes.data, n, parts, part_of.ctypes.data, load.ctypes.data)
    if rc != OK:
        raise GjdError(rc)
    return part_of[:n], load, weight[:n]


def shard_indices(files, rank, world, byte_weight=0):
    """Indices of the images rank `rank` of `world` decodes (every rank computes the same partition)."""
    part_of, load, _ = partition(files, world, byte_weight)
    return [int(i) for i in np.nonzero(part_of == rank)[0]], load
