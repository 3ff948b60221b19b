"""Parity tests proper: the CUDA path, called through the C ABI (libgjd.so), against the golden vectors
made from the unmodified reference and against the oracle on seeded synthetic inputs.
Bit-exact everywhere (coefficients, IDCT samples, pixels): integer/byte work, no tolerance."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, golden_files

pytestmark = pytest.mark.gpu
NAMES = sorted(golden_files())


@pytest.mark.parametrize("name", NAMES)
def test_coefficients_bit_exact(decoder, golden, ref_outputs, name):
    coef, w, h = decoder.decode_coeffs(golden[name])
    assert np.array_equal(coef, ref_outputs[name + "/coef"])


def test_pixels_bit_exact_batch(decoder, golden, ref_outputs):
    """Every fixture in ONE mixed-size batch (different Huffman/quant tables per image)."""
    files = [golden[n] for n in NAMES]
    outs, st = decoder.decode(files)
    assert st == [0] * len(files)
    for n, o in zip(NAMES, outs):
        assert np.array_equal(o, ref_outputs[n + "/rgb"]), n


def test_pixels_vs_reference_ground_truth(decoder, golden, ground_truth):
    for key in ground_truth.files:
        stem = key.split("/")[-1]
        outs, st = decoder.decode([golden["images/" + stem]])
        assert st == [0] and np.array_equal(outs[0], ground_truth[key]), key


def test_single_image_calls_match_batch(decoder, golden, ref_outputs):
    for n in NAMES[:6]:
        outs, st = decoder.decode([golden[n]])
        assert st == [0] and np.array_equal(outs[0], ref_outputs[n + "/rgb"])


def test_colour_exhaustive_on_device(decoder, oracle):
    """All 2^27 (Y,Cb,Cr) triples through the device colour function vs cpp-decoder/utils/color.cpp."""
    cb, cr = np.meshgrid(np.arange(-256, 256, dtype=np.int16), np.arange(-256, 256, dtype=np.int16), indexing="ij")
    cb, cr = cb.ravel(), cr.ravel()
    want = np.empty((512 * 512, 3), np.uint8)
    for y0 in range(-256, 256, 16):
        ys = np.repeat(np.arange(y0, y0 + 16, dtype=np.int16), cb.size)
        r, g, b = decoder.test_color(ys, np.tile(cb, 16), np.tile(cr, 16))
        for k in range(16):
            oracle.lib.gjdo_color_plane(y0 + k, want.ctypes.data)
            sl = slice(k * cb.size, (k + 1) * cb.size)
            assert np.array_equal(r[sl], want[:, 0]) and np.array_equal(g[sl], want[:, 1]) and np.array_equal(b[sl], want[:, 2]), y0 + k


def test_idct_on_device(decoder, oracle):
    rng = np.random.default_rng(11)
    n = 20000
    coef = np.zeros((n, 64), np.int16)
    coef[: n // 2] = rng.integers(-60, 60, (n // 2, 64)) * (rng.random((n // 2, 64)) < 0.15)
    coef[n // 2 :] = rng.integers(-1023, 1024, (n - n // 2, 64))
    coef[:, 0] = rng.integers(-2047, 2048, n)
    q = rng.integers(1, 256, 64).astype(np.uint8)
    got = decoder.test_idct(coef, q)
    zz = np.ctypeslib.as_array((__import__("ctypes").c_uint8 * 64).in_dll(oracle.lib, "gjdo_zigzag_of_natural"))
    deq = coef.astype(np.int32) * q.astype(np.int32)[None]
    nat = np.ascontiguousarray(deq[:, zz])
    for i in range(n):
        oracle.lib.gjdo_idct_block(nat[i].ctypes.data)
    assert np.array_equal(got.astype(np.int32), nat)


def test_error_statuses(decoder, golden):
    good = golden["images/3_120x120"]
    files = [good, good[: len(good) // 2], b"garbage", good]
    outs, st = decoder.decode(files)
    assert st[0] == 0 and st[3] == 0 and st[1] == -4 and st[2] == -4
    assert np.array_equal(outs[0], outs[3])
    # corrupt entropy data: flagged per image, neighbours unaffected
    data = bytearray(golden["images/5_200x200"])
    mid = len(data) // 2
    bad = bytes(data[:mid] + data[mid + 301 :])
    outs, st = decoder.decode([good, bad, good])
    assert st[0] == 0 and st[2] == 0 and st[1] in (-6, -4)
    # a frame header announcing 65535 x 65535 over a 7 KB scan: refused per image, no arena is sized from it
    at = good.find(b"\xff\xc0")
    huge = good[: at + 5] + b"\xff\xff\xff\xff" + good[at + 9 :]
    outs, st = decoder.decode([good, huge, good])
    assert st == [0, -6, 0] and outs[1] is None and np.array_equal(outs[0], outs[2])


def test_empty_batch(decoder):
    outs, st = decoder.decode([])
    assert outs == [] and st == []


def test_cli_array_output(golden, tmp_path):
    """`./decoder <file.jpg>` from a testing/-like cwd, byte-compared with the reference's committed arrays."""
    sha = json.load(open(os.path.join(GOLDEN, "array_sha256.json")))
    work = tmp_path / "testing"
    work.mkdir()
    exe = os.path.join(ROOT, "gjd-decoder", "decoder")
    for stem, want in sha.items():
        subprocess.run([exe, os.path.join(GOLDEN, "images", stem + ".jpg")], cwd=str(work), check=True)
        out = work / "gjd_output_arrays" / (stem + ".array")
        assert hashlib.sha256(out.read_bytes()).hexdigest() == want, stem
    r = subprocess.run([exe], cwd=str(work), stdout=subprocess.PIPE)
    assert r.returncode == 1 and b"Please provide the name of the image file" in r.stdout


def test_synthetic_seeded_vs_oracle(decoder, oracle):
    """Seeded synthetic images (sizes incl. non-multiples of 8, q 25..100, optimised tables) vs the oracle."""
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    specs = [(21, 1024, 1024, 90, False), (22, 777, 555, 75, True), (23, 2048, 512, 95, False),
             (24, 129, 1031, 100, False), (25, 1536, 1536, 50, True)]
    files = [gen_synth.synth_jpeg(*s) for s in specs]
    outs, st = decoder.decode(files)
    assert st == [0] * len(files)
    for f, o in zip(files, outs):
        rc, coef, rgb, _, W, H = oracle.decode(f)
        assert rc == 0 and np.array_equal(o, rgb)
        got, _, _ = decoder.decode_coeffs(f)
        assert np.array_equal(got, coef)


def test_resident_batch_and_subbatches(golden, ref_outputs):
    """Resident API, and gjd_decode_batch forced into many pipelined sub-batches by a tiny scratch hint."""
    import gjd
    files = [golden[n] for n in NAMES] * 3
    with gjd.Decoder(0, scratch_bytes=1 << 20) as d:
        outs, st = d.decode(files)
        assert st == [0] * len(files)
        for i, o in enumerate(outs):
            assert np.array_equal(o, ref_outputs[NAMES[i % len(NAMES)] + "/rgb"])
        b = d.prepare(files)
        for _ in range(2):
            b.run()
            assert b.sync() == 0
        for i in range(len(files)):
            assert np.array_equal(b.fetch(i), ref_outputs[NAMES[i % len(NAMES)] + "/rgb"])
        info = b.info()
        assert info["launches"] >= 5 and info["pixels"] > 0
        ms = d.stage_times()
        assert all(m >= 0 for m in ms)
        b.free()


def test_throughput_shape_property(decoder, oracle):
    """BASELINE config 3 shape at reduced count: 1024x1024 q90 batch with repeats — duplicates must decode
    identically (checksum of checksums), and a sample is checked against the oracle."""
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90) for s in range(8)]
    files = [base[i % 8] for i in range(96)]
    outs, st = decoder.decode(files)
    assert st == [0] * 96
    sums = [hashlib.md5(o.tobytes()).hexdigest() for o in outs]
    for i in range(96):
        assert sums[i] == sums[i % 8]
    for i in (0, 5):
        rc, _, rgb, _, W, H = oracle.decode(base[i], want_coef=False)
        assert np.array_equal(outs[i], rgb)


def test_huffman_stress_8192(decoder, oracle):
    """BASELINE config 4: one 8192x8192 q90 image without restart markers (intra-image parallel Huffman)."""
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    f = gen_synth.synth_jpeg(1000, 8192, 8192, 90)
    outs, st = decoder.decode([f])
    assert st == [0]
    rc, _, rgb, _, W, H = oracle.decode(f, want_coef=False)
    assert rc == 0 and np.array_equal(outs[0], rgb)


@pytest.mark.parametrize("sub_log2", [8, 11, 13])
def test_subsequence_lengths(golden, ref_outputs, oracle, sub_log2, monkeypatch):
    """The subsequence length is chosen per batch; force the extremes (256 bits = shorter than the look-back,
    8192 bits) and check pixels and coefficients stay bit-exact."""
    import gjd
    import gen_synth
    monkeypatch.setenv("GJD_SUB_LOG2", str(sub_log2))
    files = [golden[n] for n in NAMES]
    with gjd.Decoder(0) as d:
        outs, st = d.decode(files)
        assert st == [0] * len(files)
        for n, o in zip(NAMES, outs):
            assert np.array_equal(o, ref_outputs[n + "/rgb"]), n
        got, _, _ = d.decode_coeffs(golden[NAMES[0]])
        assert np.array_equal(got, ref_outputs[NAMES[0] + "/coef"])
        if gen_synth.have_cjpeg():
            f = gen_synth.synth_jpeg(41, 2048, 1536, 90)
            o, st = d.decode([f])
            rc, _, rgb, _, W, H = oracle.decode(f, want_coef=False)
            assert st == [0] and np.array_equal(o[0], rgb)


def test_restart_intervals_bit_exact(decoder, golden, ref_outputs, rst_golden, ref_outputs_rst):
    """DRI/RSTn (outside the reference's parser): the restart files, mixed into one batch
    with ordinary files, must give exactly what the unmodified reference gives for the twins encoded
    without restarts — coefficients and pixels."""
    rn = sorted(rst_golden)
    assert len(rn) >= 7
    for n in rn:
        coef, w, h = decoder.decode_coeffs(rst_golden[n])
        assert np.array_equal(coef, ref_outputs_rst[n + "/coef"]), n
    plain = NAMES[:5]
    files, want = [], []
    for i, n in enumerate(rn):            # interleave restart and ordinary images
        files.append(rst_golden[n]); want.append(ref_outputs_rst[n + "/rgb"])
        p = plain[i % len(plain)]
        files.append(golden[p]); want.append(ref_outputs[p + "/rgb"])
    outs, st = decoder.decode(files)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert np.array_equal(o, w_)


def test_restart_stream_errors(decoder, golden, rst_golden):
    good = golden["images/3_120x120"]
    data = bytearray(rst_golden["r23_401x363_q75_o0_rst7B"])
    i = data.index(b"\xff\xd1")
    wrong_number = bytes(data[:i + 1]) + b"\xd3" + bytes(data[i + 2:])
    j = data.index(b"\xff\xd2")
    missing = bytes(data[:j]) + bytes(data[j + 2:])
    outs, st = decoder.decode([good, wrong_number, missing, good])
    assert st[0] == 0 and st[3] == 0 and st[1] == -4 and st[2] in (-4, -6)


@pytest.mark.parametrize("sub_log2", [8, 10])
def test_restart_intervals_subsequence_lengths(rst_golden, ref_outputs_rst, sub_log2, monkeypatch):
    import gjd
    monkeypatch.setenv("GJD_SUB_LOG2", str(sub_log2))
    with gjd.Decoder(0) as d:
        rn = sorted(rst_golden)
        outs, st = d.decode([rst_golden[n] for n in rn])
        assert st == [0] * len(rn)
        for n, o in zip(rn, outs):
            assert np.array_equal(o, ref_outputs_rst[n + "/rgb"]), n


def test_throughput_driver_cli():
    """gjd-decoder/gjd_bench = the reference's benchmark_thoughput driver over the C ABI: every fixture
    decodes, in batches smaller than the directory, resident and end-to-end."""
    exe = os.path.join(ROOT, "gjd-decoder", "gjd_bench")
    assert os.path.exists(exe), "build it with make -C gpu-jpeg-decoder_b200"
    out = subprocess.run([exe, os.path.join(GOLDEN, "images"), "--batch", "4", "--iters", "2", "--e2e"],
                         check=True, stdout=subprocess.PIPE, text=True).stdout
    r = json.loads(out.strip().splitlines()[-1])
    assert r["images"] == 7 and r["failed"] == 0 and r["rejected"] == 0
    assert r["MPixel_per_s"] > 0 and r["e2e_MPixel_per_s"] > 0 and r["compressed_MB_per_s"] > 0


def test_interleaved_layout_matches_planar(decoder, golden, ref_outputs, rst_golden, ref_outputs_rst):
    """GJD_LAYOUT_INTERLEAVED (the pixel order of the reference's outputChannels / a PPM body) holds the
    same pixels as the planar layout — aligned and ragged widths, restart files included."""
    import gjd
    names = NAMES
    files = [golden[n] for n in names] + [rst_golden[n] for n in sorted(rst_golden)]
    want = [ref_outputs[n + "/rgb"] for n in names] + [ref_outputs_rst[n + "/rgb"] for n in sorted(rst_golden)]
    outs, st = decoder.decode(files, layout=gjd.INTERLEAVED)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert o.shape == (w_.shape[1], w_.shape[2], 3)
        assert np.array_equal(o, w_.transpose(1, 2, 0))


def test_interleaved_resident_fetch_and_pitch(decoder, golden, ref_outputs):
    import ctypes as C
    import gjd
    name = "images/8_401x363"
    data = golden[name]
    want = ref_outputs[name + "/rgb"].transpose(1, 2, 0)
    h, w, _ = want.shape
    pitch = 3 * w + 13                                   # unaligned pitch: the byte-store path
    items, hold = decoder.make_items([data])
    buf = np.full(pitch * h, 0xAB, np.uint8)
    items[0].rgb = buf.ctypes.data
    items[0].row_pitch = pitch
    items[0].layout = gjd.INTERLEAVED
    decoder.decode_batch_raw(items, 1)
    assert items[0].status == 0
    rows = buf.reshape(h, pitch)
    assert np.array_equal(rows[:, :3 * w].reshape(h, w, 3), want)
    items[0].row_pitch = 3 * w - 1                       # too small
    decoder.decode_batch_raw(items, 1)
    assert items[0].status == gjd.E_ARG


def test_grayscale_bit_exact(decoder, golden, ref_outputs, gray_golden, ref_outputs_gray):
    """Single-component files (outside the reference's parser), alone and mixed into a batch
    with colour files, planar and interleaved: blocks = the Y blocks of the colour twin per the unmodified
    reference, pixels = those blocks through the oracle's IDCT + colour formula with Cb = Cr = 0."""
    import gjd
    gn = sorted(gray_golden)
    assert len(gn) >= 6
    for n in gn:
        coef, w, h = decoder.decode_coeffs(gray_golden[n])
        want = ref_outputs_gray[n + "/coef"]
        assert np.array_equal(coef[: want.size], want), n
    files, want = [], []
    for i, n in enumerate(gn):
        files.append(gray_golden[n]); want.append(ref_outputs_gray[n + "/rgb"])
        p = NAMES[i % len(NAMES)]
        files.append(golden[p]); want.append(ref_outputs[p + "/rgb"])
    outs, st = decoder.decode(files)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert np.array_equal(o, w_)
    outs, st = decoder.decode(files, layout=gjd.INTERLEAVED)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert np.array_equal(o, w_.transpose(1, 2, 0))


def test_mixed_sizes_batch_independence(decoder, oracle):
    """BASELINE config 5 shape at reduced count: sizes 256..2048 (one ragged), q 75..95, some with optimised
    tables, sharded by gjd_partition into two parts.  Each part decoded as one batch must equal every image
    decoded alone (no cross-image state), and a sample must equal the oracle."""
    import gen_synth
    import gjd
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    dims = [(256, 256), (512, 512), (2048, 2048), (1024, 1024), (640, 360), (1531, 1021), (2048, 1024), (384, 1536)]
    files = [gen_synth.synth_jpeg(300 + i, w, h, 75 + 5 * (i % 5), i % 3 == 0) for i, (w, h) in enumerate(dims)] * 3
    part, load, weight = gjd.partition(files, 2)
    part, load = [int(q) for q in part], [int(v) for v in load]
    assert sorted(set(part)) == [0, 1] and sum(load) == int(weight.sum())
    assert max(load) - min(load) <= int(weight.max())          # LPT: imbalance below one item
    whole = {}
    for p in (0, 1):
        idx = [i for i, q in enumerate(part) if q == p]
        outs, st = decoder.decode([files[i] for i in idx])
        assert st == [0] * len(idx)
        whole.update(dict(zip(idx, outs)))
    for i in range(len(dims)):
        alone, st = decoder.decode([files[i]])
        assert st == [0]
        for rep in range(3):
            assert np.array_equal(whole[i + rep * len(dims)], alone[0])
    for i in (2, 5):
        rc, _, rgb, _, W, H = oracle.decode(files[i], want_coef=False)
        assert rc == 0 and np.array_equal(whole[i], rgb)


def test_fuzz_differential_vs_hostsim(decoder, hostsim, golden, rst_golden, gray_golden):
    """300 damaged files (bit flips, cuts, inserts, stray markers, truncation; headers and entropy data), decoded
    in batches between two intact files: every status equals the one the same functions give when stepped on
    the CPU (tests/hostsim), files that still decode give the same pixels, and the intact neighbours stay
    bit-exact — damage never leaks across images."""
    import fuzzlib
    seeds = ([golden[n] for n in ("images/3_120x120", "images/5_200x200", "images/6_225x168")] + list(rst_golden.values())
             + list(gray_golden.values()))
    good = golden["images/3_120x120"]
    want_good, st = decoder.decode([good])
    assert st == [0]
    bad = []
    for f in fuzzlib.cases(seeds, 300, 23):
        rc, w, h = hostsim.parse(f)
        if rc != 0 or w * h <= 1 << 20:       # a damaged SOF may ask for gigapixels: not this test
            bad.append(f)
    seen = set()
    for first in range(0, len(bad), 50):
        part = bad[first : first + 50]
        outs, st = decoder.decode([good] + part + [good])
        assert st[0] == 0 and st[-1] == 0
        assert np.array_equal(outs[0], want_good[0]) and np.array_equal(outs[-1], want_good[0])
        for f, o, s in zip(part, outs[1:-1], st[1:-1]):
            rc, _, rgb, _ = hostsim.decode(f)
            assert s == rc, (s, rc)
            seen.add(rc)
            if rc == 0:
                assert np.array_equal(o, rgb)
    assert {0, -4, -6} <= seen


def test_fuzz_differential_subsampled(decoder, hostsim, golden, sub_golden):
    """The same for chroma-subsampled seeds (4:2:0 / 4:2:2 / 4:4:0, with and without restart intervals): the ring stepper for
    MCUs of 4 or 6 blocks, the checkpointed repeat decodes and the per-tile DC chains see damaged data too."""
    import fuzzlib
    seeds = [v for k, v in sorted(sub_golden.items()) if len(v) < 60000]
    assert len(seeds) >= 5
    good = golden["images/3_120x120"]
    want_good, st = decoder.decode([good])
    assert st == [0]
    bad = []
    for f in fuzzlib.cases(seeds, 200, 29):
        rc, w, h = hostsim.parse(f)
        if rc != 0 or w * h <= 1 << 20:
            bad.append(f)
    seen = set()
    for first in range(0, len(bad), 50):
        part = bad[first : first + 50]
        outs, st = decoder.decode([good] + part + [good])
        assert st[0] == 0 and st[-1] == 0
        assert np.array_equal(outs[0], want_good[0]) and np.array_equal(outs[-1], want_good[0])
        for f, o, s in zip(part, outs[1:-1], st[1:-1]):
            rc, _, rgb, _ = hostsim.decode(f)
            assert s == rc, (s, rc)
            seen.add(rc)
            if rc == 0:
                assert np.array_equal(o, rgb)
    assert {0, -6} <= seen


def test_device_output_buffers(golden, ref_outputs, gray_golden, ref_outputs_gray):
    """rgb_on_device = 1: the caller's DEVICE buffers receive the pixels (no read-back inside the call) — tight
    planar, interleaved, and a padded pitch/plane stride — with the call forced through many sub-batches so that
    all three staging slots rotate."""
    import gjd
    import torch
    names = NAMES * 2
    files = [golden[n] for n in names] + list(gray_golden.values())
    want = [ref_outputs[n + "/rgb"] for n in names] + [ref_outputs_gray[k + "/rgb"] for k in gray_golden]
    with gjd.Decoder(0, scratch_bytes=1 << 20) as d:
        for layout, pad in ((gjd.PLANAR, 0), (gjd.INTERLEAVED, 0), (gjd.PLANAR, 24), (gjd.INTERLEAVED, 40)):
            items, hold = d.make_items(files)
            bufs = []
            for i, w in enumerate(want):
                _, H, W = w.shape
                if layout == gjd.PLANAR:
                    pitch = W + pad
                    pstride = pitch * H + 8 * pad
                    t = torch.full((3 * pstride + 64,), 0xAB, dtype=torch.uint8, device="cuda")
                    items[i].plane_stride = pstride if pad else 0
                else:
                    pitch = 3 * W + pad
                    t = torch.full((pitch * H + 64,), 0xAB, dtype=torch.uint8, device="cuda")
                items[i].row_pitch = pitch if pad else 0
                items[i].rgb = t.data_ptr()
                items[i].rgb_on_device = 1
                items[i].layout = layout
                bufs.append((t, pitch, pstride if layout == gjd.PLANAR else 0))
            d.decode_batch_raw(items, len(files))
            torch.cuda.synchronize()
            for i, w in enumerate(want):
                assert items[i].status == 0
                _, H, W = w.shape
                t, pitch, pstride = bufs[i]
                h = t.cpu().numpy()
                if layout == gjd.PLANAR:
                    got = np.stack([h[c * pstride : c * pstride + pitch * H].reshape(H, pitch)[:, :W] for c in range(3)])
                    assert np.array_equal(got, w)
                    assert (h[3 * pstride :] == 0xAB).all()          # nothing written past the third plane
                else:
                    got = h[: pitch * H].reshape(H, pitch)[:, : 3 * W].reshape(H, W, 3)
                    assert np.array_equal(got, w.transpose(1, 2, 0))
                    assert (h[pitch * H :] == 0xAB).all()


@pytest.mark.gpu
def test_transform_block_tiling_all_sizes(decoder, oracle):
    """batch_stage picks the tiles per transform block from the batch size (4, 8, 16 or 32: small batches get more
    blocks).  Resident batches of 1, 96, 160 and 320 images of 1024x1024 land on each value; every image of every
    batch must equal the oracle's pixels for its file."""
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90) for s in range(4)]
    want = []
    for f in base:
        rc, _, rgb, _, W, H = oracle.decode(f, want_coef=False)
        assert rc == 0
        want.append(hashlib.md5(rgb.tobytes()).hexdigest())
    for n in (1, 96, 160, 320):
        files = [base[i % 4] for i in range(n)]
        b = decoder.prepare(files)
        b.run()
        assert b.sync() == 0
        for i in range(n):
            assert b.status(i) == 0
            assert hashlib.md5(b.fetch(i).tobytes()).hexdigest() == want[i % 4], (n, i)
        b.free()


def test_int16_interleaved_layout(decoder, golden, ref_outputs, gray_golden, ref_outputs_gray, rst_golden, ref_outputs_rst):
    """GJD_LAYOUT_INTERLEAVED_I16: element type and order of the reference's outputChannels
    (cuda-decoder/src/parser.cu:571-573) — same pixels as the planar layout, as int16; tight, with the reference's
    padded pitch (6 * 8*ceil(W/8), the geometry write() reads back at :702-724), with an unaligned pitch, resident."""
    import gjd
    files = [golden[n] for n in NAMES] + list(gray_golden.values()) + [rst_golden[n] for n in sorted(rst_golden)]
    want = ([ref_outputs[n + "/rgb"] for n in NAMES] + [ref_outputs_gray[k + "/rgb"] for k in gray_golden]
            + [ref_outputs_rst[n + "/rgb"] for n in sorted(rst_golden)])
    outs, st = decoder.decode(files, layout=gjd.INTERLEAVED_I16)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert o.dtype == np.int16 and np.array_equal(o, w_.transpose(1, 2, 0).astype(np.int16))
    for extra in (0, 10):      # the reference's padded pitch, then the same + 10 bytes (no 16-byte alignment)
        items, hold = decoder.make_items(files)
        bufs = []
        for i, w_ in enumerate(want):
            _, H, W = w_.shape
            pitch = 6 * 8 * ((W + 7) // 8) + extra
            buf = np.full(pitch * H // 2 + 8, -21846, np.int16)
            items[i].rgb = buf.ctypes.data
            items[i].row_pitch = pitch
            items[i].layout = gjd.INTERLEAVED_I16
            bufs.append((buf, pitch))
        decoder.decode_batch_raw(items, len(files))
        for i, w_ in enumerate(want):
            assert items[i].status == 0
            _, H, W = w_.shape
            buf, pitch = bufs[i]
            rows = buf[: pitch * H // 2].reshape(H, pitch // 2)
            assert np.array_equal(rows[:, : 3 * W].reshape(H, W, 3), w_.transpose(1, 2, 0).astype(np.int16))
            # (the bytes between rows are unspecified for host buffers: include/gjd.h, regions are copied whole)
            assert (buf[pitch * H // 2:] == -21846).all()     # nothing written behind the last row
    b = decoder.prepare(files[:4], layout=gjd.INTERLEAVED_I16)
    b.run()
    assert b.sync() == 0
    for i in range(4):
        assert np.array_equal(b.fetch(i), want[i].transpose(1, 2, 0).astype(np.int16))
    b.free()
    items, hold = decoder.make_items(files[:1])
    buf = np.zeros(6 * 400 * 400 + 8, np.uint8)
    items[0].rgb = buf.ctypes.data + 1                         # odd address: not an int16 buffer
    items[0].layout = gjd.INTERLEAVED_I16
    decoder.decode_batch_raw(items, 1)
    assert items[0].status == gjd.E_ARG


def _marker_cases(base):
    sos = base.index(b"\xff\xda")
    head = base[: sos + 2 + int.from_bytes(base[sos + 2: sos + 4], "big")]
    body = base[len(head):-2]
    k = body.index(b"\xff\x00")
    return [head + body + b"\xff\xd9" + b"\xff\xc4junk\xff\xd0" * 5,      # markers behind the EOI are ignored
            head + body[:k] + b"\xff\xff\xff" + body[k:] + b"\xff\xd9",       # fill bytes in front of a stuffed FF
            head + body[:50] + b"\xff\xd0" + body[50:] + b"\xff\xd9",       # RSTn without DRI      -> -5
            head + body[:50] + b"\xff\xc4" + body[50:] + b"\xff\xd9",       # stray marker          -> -4
            head + body,                                                     # no EOI                -> -4
            head + body + b"\xff",                                           # FF as the last byte   -> -4
            head + body[:7] + b"\xff\xd9" + body]                            # early EOI: decodes short -> -6


@pytest.mark.parametrize("path", ["device", "device_pinned", "host"])
def test_unstuffing_paths_agree(golden, ref_outputs, gray_golden, ref_outputs_gray, rst_golden, ref_outputs_rst, hostsim, path, monkeypatch):
    """FF00 unstuffing + marker detection on the device (k_unstuff_*; JPEG bytes pageable or page-locked and DMA'd in
    place) and on the host (GJD_DEVICE_UNSTUFF=0) give the same pixels, coefficients and statuses — every fixture,
    restart files mixed in (they stay on the host path), hand-made marker sequences, many sub-batches."""
    import gjd
    if path == "host":
        monkeypatch.setenv("GJD_DEVICE_UNSTUFF", "0")
    pinned = path == "device_pinned"
    files = [golden[n] for n in NAMES] + list(gray_golden.values()) + [rst_golden[n] for n in sorted(rst_golden)]
    want = ([ref_outputs[n + "/rgb"] for n in NAMES] + [ref_outputs_gray[k + "/rgb"] for k in gray_golden]
            + [ref_outputs_rst[n + "/rgb"] for n in sorted(rst_golden)])
    cases = _marker_cases(golden["images/3_120x120"])
    with gjd.Decoder(0, scratch_bytes=1 << 20) as d:
        outs, st = d.decode(files + cases, pinned=pinned)
        assert st[: len(files)] == [0] * len(files)
        for o, w_ in zip(outs, want):
            assert np.array_equal(o, w_)
        got = st[len(files):]
        assert got[:2] == [0, 0] and got[2] == -5 and got[3:6] == [-4, -4, -4] and got[6] == -6
        assert np.array_equal(outs[len(files)], ref_outputs["images/3_120x120/rgb"])
        assert np.array_equal(outs[len(files) + 1], ref_outputs["images/3_120x120/rgb"])
        for f, s in zip(cases, got):
            assert s == hostsim.decode(f)[0]
        for n in NAMES[:5]:
            coef, w, h = d.decode_coeffs(golden[n])
            assert np.array_equal(coef, ref_outputs[n + "/coef"])
        with pytest.raises(gjd.GjdError) as e:
            d.decode_coeffs(cases[3])
        assert e.value.status == -4
        b = d.prepare(files[:6] + cases[2:5])
        assert [b.items[i].status for i in range(9)] == [0] * 6 + [-5, -4, -4]      # known at prepare time
        b.run()
        assert b.sync() == 0
        for i in range(6):
            assert np.array_equal(b.fetch(i), want[i])
        b.free()


def test_pinned_input_large_batch(decoder, oracle):
    """Page-locked, back-to-back JPEG bytes (the DMA fast path: one merged copy per sub-batch) at the benchmark shape."""
    import gen_synth
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    base = [gen_synth.synth_jpeg(s, 1024, 1024, 90) for s in range(4)]
    files = [base[i % 4] for i in range(64)]
    outs, st = decoder.decode(files, pinned=True)
    assert st == [0] * 64
    for i in range(4):
        rc, _, rgb, _, W, H = oracle.decode(base[i], want_coef=False)
        for j in range(i, 64, 4):
            assert np.array_equal(outs[j], rgb)


def test_subsampled_bit_exact(decoder, golden, ref_outputs, sub_golden, ref_outputs_sub):
    """4:2:0 / 4:2:2 / 4:4:0 on the GPU (k_dc_sub + k_transform_sub behind the same entropy kernels): coefficients =
    jpeg_read_coefficients of the reference's vendored libjpeg, pixels = the oracle's IDCT + sample replication + the
    oracle's colour formula (tools/make_golden_sub.py) — alone, mixed into one batch with 4:4:4 files, in every layout,
    and the reference's own testing/image.jpeg, which none of its decoders can read."""
    import gjd
    names = sorted(sub_golden)
    assert len(names) >= 8 and "ref_testing_image" in names
    for n in names:
        coef, w, h = decoder.decode_coeffs(sub_golden[n])
        assert np.array_equal(coef, ref_outputs_sub[n + "/coef"]), n
    files, want = [], []
    for i, n in enumerate(names):
        files.append(sub_golden[n]); want.append(ref_outputs_sub[n + "/rgb"])
        p = NAMES[i % len(NAMES)]
        files.append(golden[p]); want.append(ref_outputs[p + "/rgb"])
    for pinned in (False, True):
        outs, st = decoder.decode(files, pinned=pinned)
        assert st == [0] * len(files)
        for o, w_ in zip(outs, want):
            assert np.array_equal(o, w_)
    outs, st = decoder.decode(files, layout=gjd.INTERLEAVED)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert np.array_equal(o, w_.transpose(1, 2, 0))
    outs, st = decoder.decode(files, layout=gjd.INTERLEAVED_I16)
    assert st == [0] * len(files)
    for o, w_ in zip(outs, want):
        assert np.array_equal(o, w_.transpose(1, 2, 0).astype(np.int16))
    for n in names[:3]:
        outs, st = decoder.decode([sub_golden[n]])
        assert st == [0] and np.array_equal(outs[0], ref_outputs_sub[n + "/rgb"])


@pytest.mark.parametrize("sub_log2", [8, 13])
def test_subsampled_subsequence_lengths(sub_golden, ref_outputs_sub, sub_log2, monkeypatch):
    import gjd
    monkeypatch.setenv("GJD_SUB_LOG2", str(sub_log2))
    monkeypatch.setenv("GJD_DEVICE_UNSTUFF", "0" if sub_log2 == 13 else "1")
    with gjd.Decoder(0) as d:
        names = sorted(sub_golden)
        outs, st = d.decode([sub_golden[n] for n in names])
        assert st == [0] * len(names)
        for n, o in zip(names, outs):
            assert np.array_equal(o, ref_outputs_sub[n + "/rgb"]), n


def test_subsampled_large_vs_hostsim(decoder, hostsim):
    """A 4:2:0 file big enough for many sequences and transform blocks (2048x1536), against the CPU stepping of the same
    functions (itself pinned to the libjpeg/oracle composition by tests/test_hostsim.py::test_subsampled_bit_exact)."""
    import gen_synth
    import subprocess
    if not gen_synth.have_cjpeg():
        pytest.skip("vendored cjpeg missing")
    img = gen_synth.make_image(77, 2048, 1536)
    ppm = b"P6\n2048 1536\n255\n" + img.tobytes()
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(gen_synth.LIBJPEG, "lib"))
    data = subprocess.run([os.path.join(gen_synth.LIBJPEG, "bin", "cjpeg"), "-quality", "90", "-sample", "2x2", "-baseline"],
                          input=ppm, stdout=subprocess.PIPE, check=True, env=env).stdout
    outs, st = decoder.decode([data, data])
    assert st == [0, 0]
    rc, _, rgb, _ = hostsim.decode(data)
    assert rc == 0 and np.array_equal(outs[0], rgb) and np.array_equal(outs[1], rgb)


@pytest.mark.parametrize("sub_log2", [0, 12])
def test_subsampled_bench_sets_vs_pin(oracle, sub_log2, monkeypatch):
    """The 4:2:2 / 4:2:0 sets of tools/config_bench.py (1024x1024 q90: long scans, where half of the cold decodes of a 4:2:0
    subsequence start in the wrong block of the MCU and are repeated up to a checkpoint) against the pin itself, composed
    at run time as tools/make_golden_sub.py does: jpeg_read_coefficients of the vendored libjpeg + the oracle's IDCT,
    sample replication and the oracle's colour function."""
    import ctypes as C
    import gen_synth
    import gjd
    import make_golden_sub as mgs
    if not gen_synth.have_cjpeg() or not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "jpeg_coefs")):
        pytest.skip("vendored cjpeg / jpeg_coefs missing")
    if sub_log2:
        monkeypatch.setenv("GJD_SUB_LOG2", str(sub_log2))
    orc = oracle.lib
    orc.gjdo_idct_block.argtypes = [C.c_void_p]
    orc.gjdo_color_pixels.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    zz_of_nat = np.ctypeslib.as_array((C.c_uint8 * 64).in_dll(orc, "gjdo_zigzag_of_natural")).astype(np.int64)
    files, want = [], []
    for sample in ("2x1", "2x2"):
        for seed in range(3):
            data = gen_synth.synth_jpeg(seed, 1024, 1024, 90, False, 0, sample)
            coef, rgb, _ = mgs.compose(data, orc, zz_of_nat)
            files.append(data); want.append((coef, rgb))
    # restart intervals that straddle the 96-line tiles of the DC kernel: one MCU row (64 MCUs), and 11 MCUs
    for sample, restart in (("2x2", 1), ("2x1", "11B")):
        data = gen_synth.synth_jpeg(7, 1024, 1024, 90, False, restart, sample)
        coef, rgb, _ = mgs.compose(data, orc, zz_of_nat)
        files.append(data); want.append((coef, rgb))
    with gjd.Decoder(0) as d:
        for i in (0, 3, 6, 7):
            got, _, _ = d.decode_coeffs(files[i])
            assert np.array_equal(got, want[i][0])
        batch = [files[i % len(files)] for i in range(32)]
        outs, st = d.decode(batch)
        assert st == [0] * len(batch)
        for i, o in enumerate(outs):
            assert np.array_equal(o, want[i % len(files)][1]), i
