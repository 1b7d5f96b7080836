"""lsb_mod_raise_static (include/lsb200.h): the ModRaise entry point that allocates nothing -- its scratch is a file-scope
static device array of LSB_STATIC_SCRATCH_WORDS words (lattisense_b200/csrc/lsb_noalloc.cu).  It has no counterpart in the
reference, so the check is against the existing entry point lsb_mod_raise,
bit for bit as everywhere else in this suite, on two calls in a row that share the array.

Collected last on purpose (the file name sorts after every other test file).
"""
import numpy as np
import pytest

import fhe_ref as R

pytestmark = pytest.mark.gpu


def _ctx(params):
    import lattisense_b200 as L
    return L, L.Context(params.scheme, params.logN, params.q, params.p, params.t)


def _level0(params, rng, batch, npoly=2):
    return rng.integers(0, params.q[0], size=(batch, npoly, 1, 1 << params.logN), dtype=np.uint64)


def _static(L, ctx, ct0, top):
    """one call through the C ABI on buffers the caller owns (allocated before the call, as a no-malloc caller would)"""
    B, npoly, _, N = ct0.shape
    d_in = ctx.upload(ct0)
    d_out = L.DevBuf(ctx, B * npoly * (top + 1) * N)
    try:
        ctx.mod_raise_static_dev(top, npoly, d_in, d_out, B)
        return ctx.download(d_out, (B, npoly, top + 1, N))
    finally:
        d_in.free()
        d_out.free()


@pytest.mark.parametrize("params", [R.CKKS_N13, R.CKKS_BTP_N13], ids=lambda p: p.name)
def test_static_scratch_mod_raise_equals_mod_raise_twice_in_a_row(params):
    L, ctx = _ctx(params)
    rng = np.random.default_rng(77)
    top = len(params.q) - 1
    # second call: other data and a smaller batch, so anything left in the array by the first call would show
    for batch in (5, 2):
        ct0 = _level0(params, rng, batch)
        want = ctx.mod_raise(ct0, top)
        got = _static(L, ctx, ct0, top)
        assert got.shape == want.shape == (batch, 2, top + 1, 1 << params.logN)
        assert np.array_equal(got, want)


def test_static_scratch_call_allocates_no_device_memory():
    import torch

    params = R.CKKS_N13
    L, ctx = _ctx(params)
    rng = np.random.default_rng(78)
    top, N, B = len(params.q) - 1, 1 << params.logN, 4
    ct0 = _level0(params, rng, B)
    want = ctx.mod_raise(ct0, top)
    d_in = ctx.upload(ct0)
    d_out = L.DevBuf(ctx, B * 2 * (top + 1) * N)
    ctx.mod_raise_static_dev(top, 2, d_in, d_out, B)  # first use: modules load, the array's address is resolved
    ctx.sync()
    free0 = torch.cuda.mem_get_info(0)[0]
    ctx.mod_raise_static_dev(top, 2, d_in, d_out, B)
    ctx.sync()
    free1 = torch.cuda.mem_get_info(0)[0]
    assert free1 == free0
    assert np.array_equal(ctx.download(d_out, want.shape), want)
    d_in.free()
    d_out.free()


def test_static_scratch_rejects_sizes_beyond_the_array():
    params = R.CKKS_N13
    L, ctx = _ctx(params)
    N, top = 1 << params.logN, len(params.q) - 1
    B = L.STATIC_SCRATCH_WORDS // (2 * N) + 1  # one ciphertext too many
    d_in = L.DevBuf(ctx, B * 2 * N)
    d_out = L.DevBuf(ctx, B * 2 * (top + 1) * N)
    with pytest.raises(L.LsbError, match="LSB_STATIC_SCRATCH_WORDS"):
        ctx.mod_raise_static_dev(top, 2, d_in, d_out, B)
    # the largest batch that fits still runs and matches
    B -= 1
    ct0 = _level0(params, np.random.default_rng(79), B)
    ctx.upload(ct0, d_in)
    ctx.mod_raise_static_dev(top, 2, d_in, d_out, B)
    assert np.array_equal(ctx.download(d_out, (B, 2, top + 1, N)), ctx.mod_raise(ct0, top))
    d_in.free()
    d_out.free()
