"""gjd_idct_blocks_static (include/gjd.h): the allocation-free variant of gjd_test_idct, whose device scratch is a static
device array inside libgjd.so.  Checked against gjd_test_idct, bit for bit (integer work: no tolerance, as everywhere in
tests/), on two calls in a row that share the array.  This file is collected last on purpose (pytest -x)."""
import numpy as np
import pytest

import gjd


def blocks(seed, n):
    rng = np.random.default_rng(seed)
    coef = np.zeros((n, 64), np.int16)
    coef[: n // 2] = rng.integers(-60, 60, (n // 2, 64)) * (rng.random((n // 2, 64)) < 0.15)
    coef[n // 2:] = rng.integers(-1023, 1024, (n - n // 2, 64))
    coef[:, 0] = rng.integers(-2047, 2048, n)
    return coef, rng.integers(1, 256, 64).astype(np.uint8)


def test_size_beyond_static_scratch_is_rejected():
    """Refused before any CUDA call (so this one needs no GPU): GJD_E_ARG and a message that names the limit."""
    coef = np.zeros((gjd.STATIC_IDCT_MAX_BLOCKS + 1, 64), np.int16)
    with pytest.raises(gjd.GjdError) as e:
        gjd.idct_blocks_static(coef, np.ones(64, np.uint8))
    assert e.value.status == -1 and "GJD_STATIC_IDCT_MAX_BLOCKS = %d" % gjd.STATIC_IDCT_MAX_BLOCKS in str(e.value)
    assert gjd.idct_blocks_static(np.zeros((0, 64), np.int16), np.ones(64, np.uint8)).shape == (0, 64)


@pytest.mark.gpu
def test_static_scratch_matches_gjd_test_idct_two_calls_in_a_row(decoder):
    # the full array first, then a shorter call with another table: what the first call left behind must not show
    first, q1 = blocks(21, gjd.STATIC_IDCT_MAX_BLOCKS)
    second, q2 = blocks(22, 1000)
    third, q3 = blocks(23, 97)          # not a multiple of the 96 threads of a block
    want = [decoder.test_idct(c, q) for c, q in ((first, q1), (second, q2), (third, q3))]
    got1 = gjd.idct_blocks_static(first, q1)
    got2 = gjd.idct_blocks_static(second, q2)
    got3 = gjd.idct_blocks_static(third, q3)
    assert np.array_equal(got1, want[0])
    assert np.array_equal(got2, want[1])
    assert np.array_equal(got3, want[2])
    assert np.array_equal(gjd.idct_blocks_static(first, q1), want[0])      # and back again


@pytest.mark.gpu
def test_static_scratch_device_pointers_and_no_allocation(decoder):
    """The caller's buffers may be device memory; the device's free memory does not move across a call."""
    import torch
    coef, q = blocks(24, 4096)
    want = decoder.test_idct(coef, q)
    d_in = torch.from_numpy(coef).cuda()
    d_out = torch.zeros_like(d_in)
    gjd.idct_blocks_static((d_in.data_ptr(), coef.shape[0]), q, out=d_out.data_ptr())       # first call: one-time set-up
    torch.cuda.synchronize()
    d_out.zero_()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    gjd.idct_blocks_static((d_in.data_ptr(), coef.shape[0]), q, out=d_out.data_ptr())
    free_after = torch.cuda.mem_get_info(0)[0]
    assert free_after >= free_before
    assert np.array_equal(d_out.cpu().numpy(), want)
