"""CPU checks of the product's own arithmetic and algorithm (no GPU): the __host__ __device__
functions of gpu-jpeg-decoder_b200/csrc/gjd_devfn.cuh and the host parser, stepped by tests/hostsim in the
order the kernels run them, against the oracle and the golden vectors."""
import glob
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_files

NAMES = sorted(golden_files())


@pytest.mark.parametrize("name", NAMES)
def test_hostsim_bit_exact(hostsim, golden, ref_outputs, name):
    # sub_log2: the default and, to multiply the boundary cases, subsequences shorter than the look-back
    for sub_log2 in (0, 8, 12):
        rc, coef, rgb, stats = hostsim.decode(golden[name], sub_log2)
        assert rc == 0
        assert np.array_equal(coef, ref_outputs[name + "/coef"])
        assert np.array_equal(rgb, ref_outputs[name + "/rgb"])
        assert stats[2] == 0 and stats[3] == 0


def test_colour_exhaustive(hostsim, oracle):
    """gjd_color (integer fast path + IEEE fallback) == cpp-decoder/utils/color.cpp on all 2^27 triples."""
    a = np.empty((512, 512, 3), np.uint8)
    b = np.empty((512, 512, 3), np.uint8)
    for y in range(-256, 256):
        hostsim.lib.hostsim_color_plane(y, a.ctypes.data)
        oracle.lib.gjdo_color_plane(y, b.ctypes.data)
        assert np.array_equal(a, b), f"colour mismatch at y={y}"


def test_colour_exhaustive_packed_rows(hostsim, oracle):
    """gjd_color_row8 — the packed form k_transform runs (R/B in 2^-16 fixed point on 16-bit lanes, G in 2^-32 fixed
    point with the one-sided danger zone) == cpp-decoder/utils/color.cpp on all 2^27 triples."""
    a = np.empty((512, 512, 3), np.uint8)
    b = np.empty((512, 512, 3), np.uint8)
    for y in range(-256, 256):
        hostsim.lib.hostsim_color_plane_row8(y, a.ctypes.data)
        oracle.lib.gjdo_color_plane(y, b.ctypes.data)
        assert np.array_equal(a, b), f"colour mismatch at y={y}"


def test_idct_random_blocks(hostsim, oracle):
    rng = np.random.default_rng(7)
    n = 4000
    blocks = np.zeros((n, 64), np.int32)
    # sparse, dense, and extreme-range dequantised coefficients
    blocks[: n // 2] = rng.integers(-300, 300, (n // 2, 64)) * (rng.random((n // 2, 64)) < 0.2)
    blocks[n // 2 : 3 * n // 4] = rng.integers(-2048, 2048, (n // 4, 64))
    blocks[3 * n // 4 :] = rng.integers(-32768, 32768, (n - 3 * n // 4, 64))
    blocks[0] = 0
    blocks[1] = 0
    blocks[1, 0] = 1016
    out = np.zeros_like(blocks)
    hostsim.lib.hostsim_idct(blocks.ctypes.data, out.ctypes.data, n)
    want = blocks.copy()
    for i in range(n):
        oracle.lib.gjdo_idct_block(want[i].ctypes.data)
    assert np.array_equal(out, want)
    # the form k_transform runs: rows 0/4 handed over pre-multiplied, two samples finished per packed operation
    out2 = np.zeros_like(blocks)
    hostsim.lib.hostsim_idct_packed(blocks.ctypes.data, out2.ctypes.data, n)
    assert np.array_equal(out2, want)


def test_parser_contract(hostsim, golden):
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg not built (tools/_build/libjpeg)")
    import subprocess, os
    img = gen_synth.make_image(3, 64, 48)
    ppm = b"P6\n64 48\n255\n" + img.tobytes()
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(gen_synth.LIBJPEG, "lib"))
    cj = os.path.join(gen_synth.LIBJPEG, "bin", "cjpeg")

    def enc(*args):
        return subprocess.run([cj, *args], input=ppm, stdout=subprocess.PIPE, check=True, env=env).stdout

    ok = enc("-quality", "90", "-sample", "1x1", "-baseline")
    assert hostsim.parse(ok)[0] == 0
    assert hostsim.parse(enc("-sample", "2x2"))[0] == 0           # 4:2:0, 4:2:2, 4:4:0 are decoded
    assert hostsim.parse(enc("-sample", "2x1"))[0] == 0
    assert hostsim.parse(enc("-sample", "1x2"))[0] == 0
    assert hostsim.parse(enc("-sample", "4x1"))[0] == -5          # luma factors beyond 2
    assert hostsim.parse(enc("-sample", "2x2,2x1,1x1"))[0] == -5  # subsampling that leaves chroma above 1x1
    assert hostsim.parse(enc("-sample", "1x1", "-progressive"))[0] == -5
    assert hostsim.parse(enc("-grayscale"))[0] == 0                 # single component accepted
    assert hostsim.parse(enc("-sample", "1x1", "-restart", "1"))[0] == 0    # DRI/RSTn accepted
    assert hostsim.parse(enc("-sample", "1x1", "-arithmetic"))[0] == -5
    assert hostsim.parse(ok[:100])[0] == -4                       # truncated header
    assert hostsim.parse(b"not a jpeg at all")[0] == -4
    rc, *_ = hostsim.decode(ok[: len(ok) - 40])                   # truncated scan, no EOI
    assert rc == -4
    # APPn/COM segments anywhere are skipped by length (the reference parser would desynchronise)
    com = ok[:2] + b"\xff\xfe\x00\x10" + b"x" * 14 + b"\xff\xe1\x00\x08" + b"Exif\x00\x00" + ok[2:]
    rc, coef, rgb, _ = hostsim.decode(com)
    rc0, coef0, rgb0, _ = hostsim.decode(ok)
    assert rc == 0 and rc0 == 0 and np.array_equal(rgb, rgb0)


def test_corrupt_scan_is_flagged(hostsim, golden):
    data = bytearray(golden["images/5_200x200"])
    # cut a chunk out of the middle of the entropy-coded data: counts cannot add up
    mid = len(data) // 2
    bad = bytes(data[:mid] + data[mid + 301 :])
    rc, coef, rgb, stats = hostsim.decode(bad)
    assert rc in (-6, -4)


def test_hostsim_seeded_synthetic(hostsim, oracle):
    """Freshly encoded images (not in the golden set): odd sizes, q 25..100, optimised tables — the
    hostsim pipeline vs the oracle, pixels and coefficients."""
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg not built (tools/_build/libjpeg)")
    specs = [(22, 777, 555, 75, True), (24, 129, 1031, 100, False), (31, 640, 360, 25, True), (32, 1024, 1024, 90, False),
             (33, 8, 2048, 95, True), (34, 2048, 8, 50, False)]
    for sp in specs:
        data = gen_synth.synth_jpeg(*sp)
        rc, coef, rgb, stats = hostsim.decode(data)
        assert rc == 0 and stats[2] == 0, (sp, rc, list(stats))
        orc, want_coef, want_rgb, _, _, _ = oracle.decode(data)
        assert orc == 0
        assert np.array_equal(rgb, want_rgb), sp
        assert np.array_equal(coef, want_coef), sp


def test_two_level_lut_matches_canonical_tables(hostsim, golden):
    """All 65536 windows x 6 ring tables of every fixture (standard and optimised DHTs): the primary +
    secondary LUT lookup equals the canonical-table walk, and no assigned code is left to the slow path."""
    import ctypes as C
    long_codes = 0
    for name, data in sorted(golden.items()):
        out = np.zeros(3, np.uint32)
        rc = hostsim.lib.hostsim_lut_check(C.cast(C.c_char_p(data), C.c_void_p), len(data), out.ctypes.data)
        assert rc == 0, name
        assert out[0] == 0, (name, out)
        if int(os.environ.get("GJD_LUT_BITS", "10")) == 10:   # a narrower index leaves 16-bit codes to the slow path
            assert out[1] == 0, (name, out)
        long_codes += int(out[2])
    assert long_codes > 0


def test_restart_intervals_match_reference_on_twin(hostsim, rst_golden, ref_outputs_rst):
    """DRI/RSTn files (outside the reference's parser): coefficients and pixels must equal what the
    unmodified reference gives for the same image encoded without restarts (tools/make_golden_rst.py)."""
    assert len(rst_golden) >= 7
    for name, data in sorted(rst_golden.items()):
        for sub_log2 in (0, 8):
            rc, coef, rgb, stats = hostsim.decode(data, sub_log2)
            assert rc == 0 and stats[2] == 0, (name, rc, stats)
            assert np.array_equal(coef, ref_outputs_rst[name + "/coef"]), name
            assert np.array_equal(rgb, ref_outputs_rst[name + "/rgb"]), name


def test_restart_stream_errors(hostsim, rst_golden):
    data = bytearray(rst_golden["r23_401x363_q75_o0_rst7B"])
    i = data.index(b"\xff\xd1")
    bad = bytes(data[:i + 1]) + b"\xd3" + bytes(data[i + 2:])      # RST1 -> RST3: out of sequence
    assert hostsim.decode(bad)[0] == -4
    j = data.index(b"\xff\xd2")
    cut = bytes(data[:j]) + bytes(data[j + 2:])                    # one marker removed: interval count is off
    assert hostsim.decode(cut)[0] in (-4, -6)


def test_pair_counting_equals_single_symbol_counting(hostsim, golden, rst_golden, sub_golden, gray_golden):
    """k_huff_sync counts with the pair LUT (two AC symbols per trip); every span — from the true state and
    from a cold start, three span lengths, every fixture incl. q100 (blocks ending exactly at index 63) and
    restart files — must leave the decoder in the state the single-symbol walk leaves it in."""
    import ctypes as C
    trips = syms = 0
    files = dict(golden)
    files.update(rst_golden)          # restart intervals inside the scan: both walks make the same transitions
    files.update(sub_golden)          # 4:2:0 / 4:2:2 / 4:4:0 ring, with and without restarts
    files.update(gray_golden)
    for name, data in sorted(files.items()):
        for sub_log2, lookback in ((8, 100), (10, 512), (12, 512)):
            out = np.zeros(4, np.uint64)
            rc = hostsim.lib.hostsim_pair_check(C.cast(C.c_char_p(data), C.c_void_p), len(data), sub_log2, lookback,
                                                out.ctypes.data)
            assert rc == 0, name
            assert out[0] > 0 and out[1] == 0, (name, sub_log2, list(out))
            if sub_log2 == 12:
                syms += int(out[2]); trips += int(out[3])
    assert trips < 0.9 * syms        # the pairs do shorten the loop


def test_grayscale_matches_reference_pieces(hostsim, gray_golden, ref_outputs_gray):
    """Single-component files (outside the reference's parser): blocks equal the Y blocks the unmodified
    reference reports for the colour twin, pixels equal those blocks through the oracle's IDCT and colour
    formula with Cb = Cr = 0 (tools/make_golden_gray.py).  Restart files included."""
    assert len(gray_golden) >= 6
    for name, data in sorted(gray_golden.items()):
        for sub_log2 in (0, 8):
            rc, coef, rgb, stats = hostsim.decode(data, sub_log2)
            assert rc == 0 and stats[2] == 0, (name, rc, list(stats))
            want = ref_outputs_gray[name + "/coef"]
            assert np.array_equal(coef[: want.size], want), name
            assert np.array_equal(rgb, ref_outputs_gray[name + "/rgb"]), name


def _fuzz_seed_files(golden, rst_golden, gray_golden):
    return ([golden[n] for n in ("images/3_120x120", "images/5_200x200", "images/6_225x168")] + list(rst_golden.values())
            + list(gray_golden.values()))


def test_fuzz_under_address_sanitizer(tmp_path):
    """Mutation fuzzer over the parser and the device functions stepped on the CPU, built with
    -fsanitize=address,undefined: a damaged file ends in a status code, never in an out-of-bounds access, an
    out-of-range shift or index, a misaligned access, or a hang."""
    d = os.path.join(ROOT, "tests", "hostsim")
    exe = str(tmp_path / "fuzz_main")
    r = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-ffp-contract=off", "-fsanitize=address,undefined",
                        # the IDCT keeps the reference's int arithmetic (idct.cpp:24-133), which wraps on coefficients no
                        # encoder produces (same result on the device, in the oracle and here): not a finding
                        "-fno-sanitize=signed-integer-overflow", "-fno-sanitize-recover=undefined", "-Wno-unknown-pragmas",
                        "-o", exe, os.path.join(d, "fuzz_main.cpp"), os.path.join(d, "hostsim.cpp"),
                        os.path.join(ROOT, "gpu-jpeg-decoder_b200", "csrc", "gjd_parse.cpp")], capture_output=True, text=True)
    if r.returncode != 0 and ("asan" in r.stderr.lower() or "ubsan" in r.stderr.lower()):
        pytest.skip("libasan / libubsan not installed")
    assert r.returncode == 0, r.stderr[-2000:]
    files = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "rst", "*.jpg")) +
                   glob.glob(os.path.join(ROOT, "tests", "golden", "gray", "*.jpg")))
    files += [os.path.join(ROOT, "tests", "golden", "images", n) for n in ("3_120x120.jpg", "5_200x200.jpg")]
    files += [f for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "sub", "*.jpg"))) if os.path.getsize(f) < 60000]
    r = subprocess.run([exe, "3000", "7", *files], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    words = r.stdout.split()
    got = {words[i]: int(words[i + 1]) for i in range(0, len(words), 2)}
    assert got["cases"] == 3000 and got["ok"] > 0 and got["corrupt"] > 0 and got["header"] > 0


def test_fuzz_statuses_are_deterministic(hostsim, golden, rst_golden, gray_golden):
    """The same damaged file gives the same status (and, when it still decodes, the same pixels) for two
    subsequence lengths: flagged or not never depends on how the scan was cut up."""
    import fuzzlib
    n_ok = 0
    for f in fuzzlib.cases(_fuzz_seed_files(golden, rst_golden, gray_golden), 150, 11):
        rc, w, h = hostsim.parse(f)
        if rc != 0 or w * h > 1 << 20:
            continue
        a = hostsim.decode(f, 8)
        b = hostsim.decode(f, 11)
        assert a[0] == b[0] and a[0] in (0, -4, -5, -6)     # -4/-5: the unstuffer met a stray marker
        if a[0] == 0:
            n_ok += 1
            assert np.array_equal(a[2], b[2])
    assert n_ok > 5


def test_oversized_frame_header_is_refused(hostsim, golden):
    """A SOF0 that announces more 8x8 blocks than the scan bytes could hold at 2 bits per block (a damaged or
    hostile header) is refused by the parser, so it can never size a coefficient arena."""
    import gjd
    data = bytearray(golden["images/3_120x120"])
    at = data.find(b"\xff\xc0")
    assert at > 0
    big = bytes(data[: at + 5] + b"\xff\xff\xff\xff" + data[at + 9 :])     # height, width = 65535
    assert hostsim.parse(big)[0] == -6
    assert gjd.peek(big)[0] == -6
    # the largest frame the same scan could still pay for is accepted by the header check
    n_bits = (len(data) - data.find(b"\xff\xda")) * 8
    side = int((n_bits // 6) ** 0.5) * 8 // 2
    ok = bytes(data[: at + 5] + side.to_bytes(2, "big") + side.to_bytes(2, "big") + data[at + 9 :])
    assert hostsim.parse(ok)[0] == 0


def test_device_unstuffing_equals_host_loop(hostsim, golden, gray_golden, rst_golden):
    """k_unstuff_* stepped on the CPU (the product's lane functions, 16 bytes per thread, 4 KB per block) against the
    host byte loop gjd_unstuff (the twin of cpp-decoder/src/parser.cpp:78-99): same bytes, same length, and for damaged
    scans the same status — every fixture at every arena alignment, hand-made marker sequences, 600 mutated files."""
    import fuzzlib

    def has_dri(f):
        at = f.find(b"\xff\xda")
        return at >= 0 and b"\xff\xdd" in f[:at]

    files = list(golden.values()) + list(gray_golden.values()) + list(rst_golden.values())     # with and without DRI
    assert len(files) >= 30 and sum(has_dri(f) for f in files) >= 9
    for f in files:
        for mis in range(16):
            (ra, a), (rb, b) = hostsim.unstuff_both(f, mis)
            assert ra > 0 and ra == rb and a == b
    # restart markers without DRI, stray markers, fill bytes, FF at the very end, EOI followed by junk with markers
    base = golden["images/3_120x120"]
    sos = base.index(b"\xff\xda")
    head = base[: sos + 2 + int.from_bytes(base[sos + 2: sos + 4], "big")]
    body = base[len(head):-2]
    cases = [head + body + b"\xff\xd9",
             head + body[:100] + b"\xff\xff\xff" + body[100:] + b"\xff\xd9",            # fill bytes
             head + body[:4090 - len(head) % 16] + b"\xff\x00" + body[4090:] + b"\xff\xd9",   # a stuffed pair near a chunk edge
             head + body + b"\xff\xd9" + b"\xff\xc4junk\xff\xd0" * 5,                   # markers after the EOI are ignored
             head + body[:50] + b"\xff\xd0" + body[50:] + b"\xff\xd9",                    # RSTn without DRI
             head + body[:50] + b"\xff\xc4" + body[50:60] + b"\xff\xd0" + body[60:] + b"\xff\xd9",   # stray marker first
             head + body[:50] + b"\xff\xd3" + body[50:60] + b"\xff\xe1" + body[60:] + b"\xff\xd9",   # RSTn first
             head + body,                                                                # no EOI
             head + body + b"\xff",                                                      # FF as the last byte
             head + body + b"\xff\xff",
             head + b"\xff\xd9",                                                         # empty scan
             head + body[:7] + b"\xff\xd9" + body]                                       # early EOI
    for k in range(16):                                                                   # marker pairs straddling 16-byte and 4 KB edges
        cases.append(head + body[:4096 * 2 - len(head) - 8 + k] + b"\xff\x00\xff\x00\xff\xff\x00" + body[9000:] + b"\xff\xd9")
    seen = set()
    rst_seeds = [rst_golden[n] for n in sorted(rst_golden)[:3]]     # damaged restart files: wrong cycle, missing / extra markers
    for f in cases + list(fuzzlib.cases([golden[n] for n in ("images/3_120x120", "images/5_200x200", "synth/s6_256x256_q90_o1")] + rst_seeds, 900, 99)):
        for mis in (0, 5, 15):
            (ra, a), (rb, b) = hostsim.unstuff_both(f, mis)
            rc, w, h = hostsim.parse(f)
            if rc != 0:
                assert ra == rc and rb == rc
                continue
            assert ra == rb, (ra, rb, mis)
            if ra >= 0:
                assert a == b
            seen.add(ra if ra < 0 else 0)
    assert {0, -4, -5} <= seen


def test_subsampled_bit_exact(hostsim, sub_golden, ref_outputs_sub):
    """4:2:0, 4:2:2 and 4:4:0 (outside the reference's decoders): coefficients = jpeg_read_coefficients of the
    reference's vendored libjpeg, pixels = those through the oracle's IDCT + sample replication + the oracle's colour
    formula (tools/make_golden_sub.py) — ragged sizes, restart intervals, optimised tables, the reference's own
    testing/image.jpeg; default subsequences and ones shorter than the look-back."""
    assert len(sub_golden) >= 8
    for name, data in sub_golden.items():
        want_c, want_p = ref_outputs_sub[name + "/coef"], ref_outputs_sub[name + "/rgb"]
        for sub_log2 in (0, 8, 12):
            rc, coef, rgb, stats = hostsim.decode(data, sub_log2, ncoef=want_c.size)
            assert rc == 0, name
            assert np.array_equal(coef, want_c), name
            assert np.array_equal(rgb, want_p), name
            assert stats[2] == 0 and stats[3] == 0
