"""Limb-sharded key-switch (third mode) on ONE GPU: the G ranks are emulated as G sequential calls of
each phase on the same device (the kernels never wait on one another), the two all-gathers as tensor copies.  Every
limb each rank produces must equal the same limb of the oracle's single-device relinearisation."""
import numpy as np
import pytest
import torch

import fhe_ref as R
import lattisense_b200 as L
from lattisense_b200.dist import ShardedKeySwitch, shard_evk, shard_poly, unshard_ct

pytestmark = pytest.mark.gpu


def _rand_ct(o, rng, level, npoly):
    out = np.zeros((npoly, level + 1, o.N), dtype=np.uint64)
    for i in range(level + 1):
        out[:, i, :] = rng.integers(0, o.p.q[i], size=(npoly, o.N), dtype=np.uint64)
    return out


def _rand_evk(o, rng, key_level):
    shape = o.evk_shape(key_level)
    out = np.zeros(shape, dtype=np.uint64)
    for i, m in enumerate(o.p.q[:key_level + 1] + o.p.p):
        out[:, :, i, :] = rng.integers(0, m, size=(shape[0], 2, o.N), dtype=np.uint64)
    return out


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def _run_emulated(ctx, params, level, world, ct3, evk, with_addend, p2p=False):
    ranks = [ShardedKeySwitch(ctx, level, world, r, gather=lambda a, b: None) for r in range(world)]
    if p2p:  # fused all-gather: every rank's INTT kernels store into all ranks' gathered buffers (here: same GPU)
        for sk in ranks:
            sk.set_peers([x.coeff_all.data_ptr() for x in ranks], [x.accP_all.data_ptr() for x in ranks])
    c_own = [_dev(shard_poly(ct3[2], sk.own_q)) if sk.nq_own else torch.zeros((1, ctx.N), dtype=torch.int64).cuda()
             for sk in ranks]
    evk_own = [_dev(shard_evk(evk, level, params.k, sk.own_q, sk.own_p)) if sk.n_own else None for sk in ranks]
    add_own = [_dev(shard_poly(ct3[:2], sk.own_q)) if (with_addend and sk.nq_own) else None for sk in ranks]
    for sk, c in zip(ranks, c_own):
        sk.begin_p2p(c) if p2p else sk.begin(c)
    coeff_all = torch.stack([sk.coeff_own for sk in ranks])
    for sk, c, e in zip(ranks, c_own, evk_own):
        if not p2p:
            sk.coeff_all.copy_(coeff_all)
        if sk.n_own:
            sk.inner_p2p(c, e) if p2p else sk.inner(c, e)
    accp_all = torch.stack([sk.accP_own for sk in ranks])
    parts = []
    for sk, a in zip(ranks, add_own):
        if not p2p:
            sk.accP_all.copy_(accp_all)
        out = torch.zeros((2, max(sk.nq_own, 1), ctx.N), dtype=torch.int64).cuda()
        if sk.nq_own:
            sk.end(a, out)
        torch.cuda.synchronize()
        parts.append(out[:, :sk.nq_own].cpu().numpy().view(np.uint64))
    return unshard_ct(parts, level, world)


CASES = [(R.CKKS_N13, None, 2), (R.CKKS_N13, 2, 3), (R.CKKS_WIDE_N13, None, 2), (R.CKKS_WIDE_N13, 4, 4),
         (R.CKKS_N14, 3, 2), (R.CKKS_N14, None, 8), (R.CKKS_BTP_N13, 24, 8), (R.CKKS_BTP_N13, 9, 4)]


@pytest.mark.parametrize("params,level,world", CASES, ids=["%s-L%s-G%d" % (p.name, lv, w) for p, lv, w in CASES])
def test_sharded_relin_matches_oracle(params, level, world):
    o = R.Oracle(params)
    ctx = L.Context(params.scheme, params.logN, params.q, params.p, params.t)
    level = params.max_level if level is None else level
    rng = np.random.default_rng(100 + world)
    ct3 = _rand_ct(o, rng, level, 3)
    evk = _rand_evk(o, rng, level)
    want = o.relin(ct3, evk)
    got = _run_emulated(ctx, params, level, world, ct3, evk, True)
    assert np.array_equal(got, want)
    # the all-gathers fused into the INTT kernels (stores into every rank's gathered buffer)
    assert np.array_equal(_run_emulated(ctx, params, level, world, ct3, evk, True, p2p=True), want)
    # without the (c0, c1) addend: the bare key-switch of c2 (what a rotation adds c0 to)
    ct3z = ct3.copy()
    ct3z[:2] = 0
    assert np.array_equal(_run_emulated(ctx, params, level, world, ct3z, evk, False), o.relin(ct3z, evk))
    ctx.close()


def test_n65536_sharded_over_8_ranks():
    params = R.CKKS_BTP_N16
    o = R.Oracle(params)
    ctx = L.Context(params.scheme, params.logN, params.q, params.p, params.t)
    level = 24
    rng = np.random.default_rng(7)
    ct3 = _rand_ct(o, rng, level, 3)
    evk = _rand_evk(o, rng, level)
    assert np.array_equal(_run_emulated(ctx, params, level, 8, ct3, evk, True), o.relin(ct3, evk))
    ctx.close()


@pytest.mark.parametrize("params,level,batch", [(R.CKKS_N13, None, 3), (R.CKKS_BTP_N13, 24, 2)], ids=["ckks8k-B3", "btp8k-L24-B2"])
def test_cpp_nccl_driver_world1_matches_oracle(params, level, batch):
    """The C++ host driver (lsb_ks_shard_create / lsb_ks_shard_run: NCCL loaded at run time, communicator, exchange
    buffers, CUDA graph) on a one-rank world: the all-gathers degenerate to copies, every limb must equal the oracle's
    relinearisation; the eager first call, the capturing second call and the graph replay give the same limbs."""
    from lattisense_b200.dist import NcclShardedKeySwitch
    o = R.Oracle(params)
    ctx = L.Context(params.scheme, params.logN, params.q, params.p, params.t)
    level = params.max_level if level is None else level
    rng = np.random.default_rng(31)
    cts = [_rand_ct(o, rng, level, 3) for _ in range(batch)]
    evk = _rand_evk(o, rng, level)
    sk = NcclShardedKeySwitch(ctx, level, 1, 0, max_batch=batch)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        c_own = _dev(np.stack([ct[2] for ct in cts]))
        add_own = _dev(np.stack([ct[:2] for ct in cts]))
        evk_d = _dev(evk)
        out = torch.empty((batch, 2, level + 1, ctx.N), dtype=torch.int64, device="cuda")
        want = np.stack([o.relin(ct, evk) for ct in cts])
        for it in range(3):  # eager, capture + launch, replay
            out.zero_()
            sk.run(c_own, evk_d, add_own, out=out, stream=s.cuda_stream)
            s.synchronize()
            assert np.array_equal(out.cpu().numpy().view(np.uint64), want), "iteration %d" % it
    assert sk.all_gather_bytes() == batch * ((level + 1) + 2 * len(params.p)) * ctx.N * 8
    sk.close()
    ctx.close()
