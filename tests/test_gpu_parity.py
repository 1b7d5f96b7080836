"""GPU parity tests: every output limb of the CUDA path (called through the C ABI in
include/lsb200.h) must equal the CPU oracle bit for bit on the same seeded inputs.

Reference test cases mirrored (unittests/test_gpu_bfv.cpp / test_gpu_ckks.cpp, decode-level there):
cmc, cmc_relin (:314), cmc_relin_rescale (:340), rescale, drop_level, rotate_col (:459),
advanced_rotate_col (:493), rotate_row (:530), cac/csc/cneg, cap/csp, cmp.  The reference compares
decoded messages only; limb equality against the oracle is the stricter bar set by BASELINE.json.
"""
import numpy as np
import pytest

import fhe_ref as R

pytestmark = pytest.mark.gpu


def _ctx(params):
    import lattisense_b200 as L
    return L.Context(params.scheme, params.logN, params.q, params.p, params.t)


_cache = {}


def env(params):
    """(oracle, gpu context, secret) per parameter set, cached for the session."""
    if params.name not in _cache:
        o = R.Oracle(params)
        _cache[params.name] = (o, _ctx(params), o.gen_secret(1234))
    return _cache[params.name]


def rand_ct(o, rng, level, npoly=2, batch=None):
    shape = (npoly, level + 1, o.N) if batch is None else (batch, npoly, level + 1, o.N)
    out = np.zeros(shape, dtype=np.uint64)
    for i in range(level + 1):
        out[..., i, :] = rng.integers(0, o.p.q[i], size=shape[:-2] + (o.N,), dtype=np.uint64)
    return out


def rand_evk(o, rng, key_level):
    shape = o.evk_shape(key_level)
    out = np.zeros(shape, dtype=np.uint64)
    mods = o.p.q[:key_level + 1] + o.p.p
    for i, m in enumerate(mods):
        out[:, :, i, :] = rng.integers(0, m, size=(shape[0], 2, o.N), dtype=np.uint64)
    return out


SETS = [R.CKKS_N12, R.CKKS_N13, R.CKKS_N14, R.BFV_N13, R.BFV_N14, R.CKKS_WIDE_N13]
IDS = [p.name for p in SETS]


@pytest.mark.parametrize("params", SETS + [R.CKKS_BTP_N15, R.CKKS_BTP_N16], ids=IDS + ["btp_n32768", "btp_n65536"])
def test_ntt_parity_all_moduli(params):
    o, g, _ = env(params)
    rng = np.random.default_rng(21)
    nm = o.nmods if params.logN <= 14 else 8
    mids = list(range(nm)) if params.logN <= 14 else [0, 1, 9, 12, 24, 25, 27, 29]
    a = np.stack([rng.integers(0, o.moduli[m], size=o.N, dtype=np.uint64) for m in mids])
    a[:, :3] = 0
    a[0, :3] = [1, o.moduli[mids[0]] - 1, 2]
    batch = np.stack([a, a[:, ::-1].copy(), np.roll(a, 17, axis=1)])
    fwd = g.ntt(batch, mids)
    inv = g.ntt(batch, mids, inverse=True)
    for b in range(3):
        for r, m in enumerate(mids):
            assert np.array_equal(fwd[b, r], o.ntt(m, batch[b, r])), (b, m)
            assert np.array_equal(inv[b, r], o.ntt(m, batch[b, r], inverse=True)), (b, m)
    assert np.array_equal(g.ntt(fwd, mids, inverse=True), batch)


@pytest.mark.parametrize("params", SETS, ids=IDS)
def test_mult_relin_rescale_parity(params):
    o, g, s = env(params)
    rng = np.random.default_rng(31)
    L = params.max_level
    rlk = o.gen_rlk(s, 77, L)
    for level in sorted({L, max(1, L // 2), 1} | ({4, 5} if L >= 5 else set())):
        a, b = rand_ct(o, rng, level), rand_ct(o, rng, level)
        want3 = o.mult(a, b)
        got3 = g.mult(a, b)
        assert np.array_equal(got3, want3), ("mult", level)
        want2 = o.relin(want3, rlk)
        assert np.array_equal(g.relin(want3, rlk), want2), ("relin", level)
        want1 = o.rescale(want2)
        assert np.array_equal(g.rescale(want2), want1), ("rescale", level)
        assert np.array_equal(g.rescale(want3), o.rescale(want3)), ("rescale ct3", level)
        assert np.array_equal(g.mult_relin(a, b, rlk), want2), ("mult_relin", level)
        assert np.array_equal(g.mult_relin(a, b, rlk, rescale=True), want1), ("mult_relin_rescale", level)
        assert np.array_equal(g.drop_level(a), o.drop_level(a))


@pytest.mark.parametrize("params", SETS, ids=IDS)
def test_rotate_parity(params):
    o, g, s = env(params)
    rng = np.random.default_rng(41)
    L = params.max_level
    for step, level in [(1, L), (3, max(1, L - 1)), (-5, 1)]:
        gal = R.galois_for_col_step(o.N, step)
        glk = o.gen_glk(s, 90 + step, L, gal)
        ct = rand_ct(o, rng, level)
        assert np.array_equal(g.rotate(ct, glk, gal), o.rotate(ct, glk, gal)), (step, level)
    gal = R.galois_for_row(o.N)
    glk = rand_evk(o, rng, L)
    ct = rand_ct(o, rng, L)
    assert np.array_equal(g.rotate(ct, glk, gal), o.rotate(ct, glk, gal))


@pytest.mark.parametrize("params", [R.CKKS_N13, R.BFV_N13], ids=["ckks", "bfv"])
def test_batched_and_ragged_chunks(params):
    """batch sizes that do not divide the chunk, chunk = 1, and batch = 1 all give per-item oracle results"""
    o, g, s = env(params)
    rng = np.random.default_rng(51)
    L = params.max_level
    rlk = rand_evk(o, rng, L)
    B = 5
    a, b = rand_ct(o, rng, L, batch=B), rand_ct(o, rng, L, batch=B)
    want = np.stack([o.rescale(o.relin(o.mult(a[i], b[i]), rlk)) for i in range(B)])
    gal = R.galois_for_col_step(o.N, 2)
    want_rot = np.stack([o.rotate(a[i], rlk, gal) for i in range(B)])
    try:
        for chunk in (0, 1, 2, 3, 7):
            g.set_chunk(chunk)
            assert np.array_equal(g.mult_relin(a, b, rlk, rescale=True), want), chunk
            assert np.array_equal(g.rotate(a, rlk, gal), want_rot), chunk
    finally:
        g.set_chunk(0)


@pytest.mark.parametrize("params", SETS, ids=IDS)
def test_add_sub_neg_parity(params):
    o, g, _ = env(params)
    rng = np.random.default_rng(61)
    L = params.max_level
    a, b = rand_ct(o, rng, L, batch=2), rand_ct(o, rng, L, batch=2)
    a[0, 0, 0, :4] = 0
    b[0, 0, 0, :4] = 0
    for i in range(2):
        assert np.array_equal(g.add(a, b)[i], o.add(a[i], b[i]))
        assert np.array_equal(g.sub(a, b)[i], o.sub(a[i], b[i]))
        assert np.array_equal(g.neg(a)[i], o.neg(a[i]))
    a3 = rand_ct(o, rng, 1, npoly=3)
    assert np.array_equal(g.add(a3, a3), o.add(a3, a3))


@pytest.mark.parametrize("params", SETS, ids=IDS)
def test_plain_ops_parity(params):
    """ct +- pt, ct x pt_mul, ct (+,x) pt_ringt against the oracle's limb arithmetic"""
    o, g, _ = env(params)
    rng = np.random.default_rng(71)
    level = min(2, params.max_level)
    l = level + 1
    ct = rand_ct(o, rng, level)
    pt = rand_ct(o, rng, level, npoly=1)[0]
    want = ct.copy()
    want[0] = o.add(ct[:1], pt[None])[0]
    assert np.array_equal(g.add_plain(ct, pt), want)
    want[0] = o.sub(ct[:1], pt[None])[0]
    assert np.array_equal(g.add_plain(ct, pt, subtract=True), want)

    def mulmod_limbs(x, y):
        out = np.zeros_like(x)
        for i in range(l):
            q = params.q[i]
            out[:, i] = np.array([[int(u) * int(v) % q for u, v in zip(xr, y[i])] for xr in x[:, i]], dtype=np.uint64)
        return out

    ntt_dom = params.scheme == R.SCHEME_CKKS
    ct_n = ct if ntt_dom else o.ntt_limbs(ct)
    prod = mulmod_limbs(ct_n, pt)
    want_mul = prod if ntt_dom else o.ntt_limbs(prod, inverse=True)
    assert np.array_equal(g.mul_plain(ct, pt), want_mul)

    t = params.t if params.scheme == R.SCHEME_BFV else 65537
    m = rng.integers(0, t, size=(1, o.N), dtype=np.uint64)
    lifted = o.ntt_limbs(np.repeat(m, l, axis=0))
    prod = mulmod_limbs(ct_n, lifted)
    want_mr = prod if ntt_dom else o.ntt_limbs(prod, inverse=True)
    assert np.array_equal(g.mul_plain_ringt(ct, m), want_mr)
    if params.scheme == R.SCHEME_BFV:
        up = o.bfv_scale_up(level, m[0])
        want = ct.copy()
        want[0] = o.add(ct[:1], up[None])[0]
        assert np.array_equal(g.add_plain_ringt(ct, m), want)
    else:
        want = ct.copy()
        want[0] = o.add(ct[:1], lifted[None])[0]
        assert np.array_equal(g.add_plain_ringt(ct, m), want)


def test_bfv_mult_relin_decrypts_to_product():
    """reference criterion (unittests/test_gpu_bfv.cpp:314-335): decode(dec(z)) == x*y mod t"""
    params = R.BFV_N13
    o, g, s = env(params)
    rng = np.random.default_rng(81)
    enc = R.BfvEncoder(params)
    L = params.max_level
    x = rng.integers(0, params.t, size=o.N, dtype=np.uint64)
    y = rng.integers(0, params.t, size=o.N, dtype=np.uint64)
    cx = o.encrypt(s, 5, L, o.bfv_scale_up(L, enc.encode_coeffs(x)))
    cy = o.encrypt(s, 6, L, o.bfv_scale_up(L, enc.encode_coeffs(y)))
    rlk = o.gen_rlk(s, 7, L)
    z = g.mult_relin(cx, cy, rlk)
    got = enc.decode_coeffs(R.bfv_decode_pt(o, o.decrypt(s, z)))
    assert np.array_equal(got, x * y % params.t)
    gal = R.galois_for_col_step(o.N, 1)
    glk = o.gen_glk(s, 8, L, gal)
    r = enc.decode_coeffs(R.bfv_decode_pt(o, o.decrypt(s, g.rotate(cx, glk, gal))))
    half = o.N // 2
    want = np.concatenate([np.roll(x[:half], -1), np.roll(x[half:], -1)])
    assert np.array_equal(r, want)


def test_ckks_mult_relin_rescale_precision():
    """reference criterion (unittests/test_gpu_ckks.cpp:37-43): mean real precision >= 10 bits"""
    params = R.CKKS_N13
    o, g, s = env(params)
    rng = np.random.default_rng(91)
    L = params.max_level
    x = rng.uniform(-1, 1, o.N // 2)
    y = rng.uniform(-1, 1, o.N // 2)
    cx = o.encrypt(s, 15, L, R.ckks_encode(o, x, L, params.scale))
    cy = o.encrypt(s, 16, L, R.ckks_encode(o, y, L, params.scale))
    rlk = o.gen_rlk(s, 17, L)
    z = g.mult_relin(cx, cy, rlk, rescale=True)
    scale = params.scale * params.scale / params.q[L]
    got = R.ckks_decode(o, o.decrypt(s, z), scale)
    assert R.mean_precision_bits(got, x * y) >= 10.0
    gal = R.galois_for_col_step(o.N, 3)
    glk = o.gen_glk(s, 18, L, gal)
    r = R.ckks_decode(o, o.decrypt(s, g.rotate(cx, glk, gal)), params.scale)
    assert R.mean_precision_bits(r, np.roll(x, -3)) >= 10.0


def test_n65536_keyswitch_parity():
    """cfg4/cfg5 shapes: N = 2^16 with the 25Q+5P bootstrap chain (frontend/custom_task.py:387-420)"""
    params = R.CKKS_BTP_N16
    o, g, s = env(params)
    rng = np.random.default_rng(101)
    for level in (2, 24):
        key_level = 24 if level == 24 else 4
        evk = rand_evk(o, rng, key_level)
        a, b = rand_ct(o, rng, level), rand_ct(o, rng, level)
        ct3 = o.mult(a, b)
        assert np.array_equal(g.mult(a, b), ct3)
        want = o.relin(ct3, evk)
        assert np.array_equal(g.relin(ct3, evk), want), level
        if level == 2:
            gal = R.galois_for_col_step(o.N, 16)
            assert np.array_equal(g.rotate(a, evk, gal), o.rotate(a, evk, gal))
            assert np.array_equal(g.rescale(want), o.rescale(want))


@pytest.mark.parametrize("params,levels", [(R.CKKS_MANYDIGIT_N13, (24, 17, 13)), (R.CKKS_N15, (17, 8)), (R.CKKS_N16, (29, 27, 12))],
                         ids=["many_digits_n8192", "default_n32768", "default_n65536"])
def test_keyswitch_with_more_digits_than_one_launch_group(params, levels):
    """Key switches whose beta * (l + k) extended limbs exceed one launch's 192 rows / 256 slots are run in digit groups
    (Context::ks_digit_group): the reference frontend's default n = 65536 CKKS set at levels 27..29 (beta = 7..8 over 34
    limbs), and a 2-limb-digit chain with beta = 13 on a small ring; the default n = 32768 set rides along (one group)."""
    o, g, s = env(params)
    rng = np.random.default_rng(211)
    top = params.max_level
    evk = rand_evk(o, rng, top)
    for level in levels:
        a, b = rand_ct(o, rng, level), rand_ct(o, rng, level)
        ct3 = o.mult(a, b)
        want = o.relin(ct3, evk)
        assert np.array_equal(g.relin(ct3, evk), want), ("relin", level)
        assert np.array_equal(g.mult_relin(a, b, evk, rescale=True), o.rescale(want)), ("mult_relin_rescale", level)
        gal = R.galois_for_col_step(o.N, 3)
        assert np.array_equal(g.rotate(a, evk, gal), o.rotate(a, evk, gal)), ("rotate", level)
    if params.logN <= 13:
        # hoisted rotations share the grouped decomposition
        level = levels[0]
        a = rand_ct(o, rng, level)
        gals = [R.galois_for_col_step(o.N, st) for st in (1, 2, 5)]
        keys = [rand_evk(o, rng, top) for _ in gals]
        got = g.rotate_hoisted(a, keys, gals)
        for i, gal in enumerate(gals):
            assert np.array_equal(got[i], o.rotate(a, keys[i], gal)), ("hoisted", gal)


@pytest.mark.parametrize("params", [R.BFV_N12, R.CKKS_N12_DEFAULT, R.BFV_N15], ids=lambda p: p.name)
def test_remaining_default_parameter_sets(params):
    """frontend/parameter.json's default sets that no other test uses (n = 4096 for both schemes, BFV n = 32768 with its
    12 Q + 3 P limbs of 58-61 bits): mult / relin / rescale / rotate at the top level and at level 1"""
    o, g, s = env(params)
    rng = np.random.default_rng(223)
    top = params.max_level
    rlk = o.gen_rlk(s, 5, top)
    gal = R.galois_for_col_step(o.N, 2)
    glk = o.gen_glk(s, 6, top, gal)
    for level in sorted({top, 1}):
        a, b = rand_ct(o, rng, level), rand_ct(o, rng, level)
        want = o.relin(o.mult(a, b), rlk)
        assert np.array_equal(g.mult_relin(a, b, rlk), want), ("mult_relin", level)
        assert np.array_equal(g.rescale(want), o.rescale(want)), ("rescale", level)
        assert np.array_equal(g.rotate(a, glk, gal), o.rotate(a, glk, gal)), ("rotate", level)


def test_calls_on_two_streams_share_the_scratch_block_safely():
    """One context, its default scratch block, two streams, no synchronisation by the caller between the calls: the
    library orders a call after the previous stream's work when the stream changes (the scratch is still in use there),
    so both results equal the ones of the same calls made one after the other."""
    import torch
    params = R.CKKS_N14
    o, g, s = env(params)
    rng = np.random.default_rng(227)
    level, B = 3, 96
    rlk = rand_evk(o, rng, level)
    a, b = rand_ct(o, rng, level, batch=B), rand_ct(o, rng, level, batch=B)
    c = rand_ct(o, rng, level, batch=B)
    want_ab = g.mult_relin(a, b, rlk)
    want_ac = g.mult_relin(a, c, rlk)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    da, db, dc, dk = g.upload(a), g.upload(b), g.upload(c), g.upload(rlk)
    import lattisense_b200 as L
    o1, o2 = L.DevBuf(g, want_ab.size), L.DevBuf(g, want_ac.size)
    keep = g.stream
    try:
        for _ in range(3):
            g.stream = s1.cuda_stream
            g.mult_relin_dev(level, da, db, dk, level, o1, B)
            g.stream = s2.cuda_stream
            g.mult_relin_dev(level, da, dc, dk, level, o2, B)
        s1.synchronize()
        s2.synchronize()
    finally:
        g.stream = keep
    assert np.array_equal(g.download(o1, want_ab.shape), want_ab)
    assert np.array_equal(g.download(o2, want_ac.shape), want_ac)
    for d in (da, db, dc, dk, o1, o2):
        d.free()


def test_large_batches_split_over_two_streams_match_the_oracle():
    """batches of 256 and more run as two half-batches on two streams (mult_relin chain and rotate, LSB_DUAL_STREAM): items
    on both sides of the split, the first and the last equal the oracle's"""
    params = R.CKKS_N13
    o, g, s = env(params)
    rng = np.random.default_rng(229)
    level, B = 3, 301
    rlk = rand_evk(o, rng, level)
    gal = R.galois_for_col_step(o.N, 7)
    a, b = rand_ct(o, rng, level, batch=B), rand_ct(o, rng, level, batch=B)
    rot = g.rotate(a, rlk, gal)
    mrr = g.mult_relin(a, b, rlk, rescale=True)
    for i in (0, 150, 151, B - 1):
        assert np.array_equal(rot[i], o.rotate(a[i], rlk, gal)), ("rotate", i)
        assert np.array_equal(mrr[i], o.rescale(o.relin(o.mult(a[i], b[i]), rlk))), ("mult_relin_rescale", i)


def test_errors_are_reported_not_swallowed():
    import lattisense_b200 as L
    params = R.CKKS_N12
    o, g, _ = env(params)
    rng = np.random.default_rng(111)
    ct0 = rand_ct(o, rng, 0)
    with pytest.raises(L.LsbError):
        g.rescale(ct0)
    evk_low = rand_evk(o, rng, 1)
    ct3 = rand_ct(o, rng, 2, npoly=3)
    with pytest.raises(L.LsbError):
        g.relin(ct3, evk_low)
    with pytest.raises(L.LsbError):
        L.Context(R.SCHEME_CKKS, 12, [params.q[0] + 2], params.p)
    with pytest.raises(L.LsbError):
        L.Context(R.SCHEME_CKKS, 11, params.q, params.p)


@pytest.mark.gpu
def test_toy_bootstrap_device_ops_equal_the_oracle_limb_for_limb():
    """The whole bootstrapping pipeline of tests/btp_pipeline.py at N=2^12 (ModRaise, CoeffsToSlots, EvalMod,
    SlotsToCoeffs: 154 rotations / conjugations, 28 ct x ct + relin, 44 rescales, 55-bit primes, 15 levels deep) run
    twice on the same keys: once on the oracle, once with every ciphertext operation on the device through lsb200.h.  Output limbs must be
    identical and decode to the input within the reference's bootstrapping criterion."""
    import btp_pipeline as P
    import test_btp_model as TB

    _, R, prm, orc, secret, z, ct0 = TB._toy_case(12, 17)  # the device supports N >= 2^12
    want = P.bootstrap(P.Evaluator(orc, secret), ct0, cts_depth=3, stc_depth=3)
    ctx = _ctx(prm)
    ev = P.Evaluator(orc, secret, ops=ctx)
    got = P.bootstrap(ev, ct0, cts_depth=3, stc_depth=3)
    assert got.level == want.level == 2
    assert np.array_equal(got.data, want.data)
    assert R.mean_precision_bits(ev.decode(got, secret), z) >= 12.0


@pytest.mark.gpu
@pytest.mark.parametrize("params", [R.CKKS_N13, R.CKKS_WIDE_N13, R.CKKS_BTP_N13], ids=lambda p: p.name)
def test_mod_raise_parity(params):
    """lsb_mod_raise against the pipeline's ModRaise on the oracle (centred residues mod q0 re-read mod every prime)."""
    import btp_pipeline as P

    o, ctx, s = env(params)
    rng = np.random.default_rng(21)
    top = len(params.q) - 1
    ct0 = rand_ct(o, rng, 0, batch=3)
    ev = P.Evaluator(o, s)
    want = np.stack([P.mod_raise(ev, P.Ct(ct0[b], 1.0)).data for b in range(3)])
    got = ctx.mod_raise(ct0, top)
    assert got.shape == want.shape == (3, 2, top + 1, o.N)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_linear_transform_parity():
    """lsb_linear_transform (BSGS: rotations, one multiply-accumulate pass per giant step, rescale) against the same
    matrix evaluated op by op on the oracle; the matrix is a merged CoeffsToSlots factor (tests/btp_model.py)."""
    import btp_model as B
    import btp_pipeline as P

    prm = P.toy_params(12, 4)
    o = R.Oracle(prm)
    secret = P.sparse_secret(o.N, 32, 9)
    n = o.N // 2
    rng = np.random.default_rng(5)
    z = rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)
    ct = P.Ct(o.encrypt(secret, 4, 4, R.ckks_encode(o, z, 4, 2.0 ** 45)), 2.0 ** 45)
    mats = B.factor_c2s(n, 3)
    idx = B.dft_index_map(n.bit_length() - 1, 3, True)
    for m, ix in zip(mats[:2], idx[:2]):
        n1 = B.bsgs_split(ix, n, 2.0)
        want = P.Evaluator(o, secret).linear(ct, m, n1, float(prm.q[4]))
        ev = P.Evaluator(o, secret, ops=_ctx(prm))
        got = ev.linear(ct, m, n1, float(prm.q[4]))
        assert np.array_equal(got.data, want.data)
        assert np.abs(ev.decode(got, secret) - B.apply(m, z)).max() < 1e-6


@pytest.mark.parametrize("params", [R.CKKS_N13, R.BFV_N13], ids=lambda p: p.name)
@pytest.mark.parametrize("m,with_acc,ringt", [(2, False, False), (5, True, False), (20, False, True), (3, True, True)])
def test_mac_parity(params, m, with_acc, ringt):
    """lsb_mac (bind_gpu_cmp_sum / bind_gpu_cmpac_sum, mega_ag_executors_gpu.cu:294-408; reference cases
    unittests/test_gpu_ckks.cpp:675-705 `ct_pt_mac` m = 2..20) against the node-by-node oracle evaluation: NTT-domain
    plaintexts (CKKS pt / BFV pt_mul) and pt_ringt, with and without a partial sum, batch of 2."""
    o = R.Oracle(params)
    g = _ctx(params)
    rng = np.random.default_rng(m)
    lv = params.max_level - 1
    l = lv + 1
    bfv = params.scheme == R.SCHEME_BFV
    q = params.q[:l]
    rand = lambda lead: np.stack([rng.integers(0, mod, size=lead + (o.N,), dtype=np.uint64) for mod in q], axis=-2)
    B = 2
    cts = [rand((B, 2)) for _ in range(m)]
    if ringt:
        t = params.t if bfv else 65537
        pts = [rng.integers(0, t, size=(B, 1, o.N), dtype=np.uint64) for _ in range(m)]
        pts_ntt = [np.stack([o.ntt_limbs(np.repeat(p[b], l, axis=0)) for b in range(B)]) for p in pts]
    else:
        pts = [rand((B,)) for _ in range(m)]
        pts_ntt = pts
    acc = rand((B, 2)) if with_acc else None
    want = np.zeros((B, 2, l, o.N), dtype=np.uint64)
    for b in range(B):
        tot = None
        for i in range(m):
            x = o.ntt_limbs(cts[i][b]) if bfv else cts[i][b]
            term = np.zeros_like(x)
            for p in range(2):
                for j in range(l):
                    term[p, j] = ((x[p, j].astype(object) * pts_ntt[i][b][j].astype(object)) % q[j]).astype(np.uint64)
            tot = term if tot is None else o.add(tot, term)
        out = o.ntt_limbs(tot, inverse=True) if bfv else tot
        want[b] = o.add(out, acc[b]) if with_acc else out
    got = g.mac(cts, pts, ringt=ringt, acc=acc)
    assert np.array_equal(got, want)
    g.close()


@pytest.mark.parametrize("params,level", [(R.CKKS_N13, None), (R.CKKS_BTP_N13, 20), (R.CKKS_WIDE_N13, 4)], ids=["ckks8k", "btp8k-L20", "wide8k-L4"])
def test_hoisted_rotations_equal_plain_rotations(params, level):
    """lsb_rotate_hoisted (one decomposition shared by several rotations of the same ciphertext) gives, for every rotation,
    the limbs of the oracle's rotate -- the same as lsb_rotate."""
    o = R.Oracle(params)
    g = _ctx(params)
    level = params.max_level if level is None else level
    rng = np.random.default_rng(9)
    s = o.gen_secret(4)
    ct = np.stack([np.stack([rng.integers(0, q, size=(2, o.N), dtype=np.uint64) for q in params.q[:level + 1]], axis=1) for _ in range(2)])
    steps = [1, 5, 100, o.N // 2 - 3]
    gals = [R.galois_for_col_step(o.N, st) for st in steps] + [R.galois_for_row(o.N)]
    glks = [o.gen_glk(s, 50 + i, level, gal) for i, gal in enumerate(gals)]
    got = g.rotate_hoisted(ct, glks, gals)
    for r, (gal, glk) in enumerate(zip(gals, glks)):
        for b in range(2):
            assert np.array_equal(got[r][b], o.rotate(ct[b], glk, gal)), (r, b)
    g.close()


@pytest.mark.parametrize("params", [R.CKKS_N13, R.CKKS_N14, R.CKKS_WIDE_N13], ids=lambda p: p.name)
def test_keyswitch_overflow_estimate_near_integers(params):
    """The FP64 CRT kernel forms Lattigo's float64 overflow count with multiplications and falls back to the exact
    division sequence when the sum lies within 2^-40 of an integer.  Digits whose CRT value is 0, +-1, +-2 (sum = k + tiny or
    k - tiny) sit exactly there: the relinearisation of such a c2 must still equal the oracle's limb for limb."""
    o = R.Oracle(params)
    g = _ctx(params)
    level = params.max_level
    l = level + 1
    rng = np.random.default_rng(77)
    small = rng.integers(-2, 3, size=o.N)  # coefficient-domain c2 with values in {-2..2}: every digit's CRT value is tiny or Q_d - tiny
    small[:8] = [0, 1, -1, 2, -2, 0, 1, -1]
    c2c = np.stack([(small % q).astype(np.uint64) for q in params.q[:l]])
    ct3 = np.stack([np.stack([rng.integers(0, q, size=o.N, dtype=np.uint64) for q in params.q[:l]]) for _ in range(2)] + [o.ntt_limbs(c2c)])
    evk = np.stack([np.stack([np.stack([rng.integers(0, m, size=o.N, dtype=np.uint64) for m in params.q[:l] + params.p]) for _ in range(2)])
                    for _ in range(params.beta(level))])
    assert np.array_equal(g.relin(ct3, evk), o.relin(ct3, evk))
    g.close()
