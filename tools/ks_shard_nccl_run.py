"""Limb-sharded key-switch through the C++ NCCL driver (lsb_ks_shard_create / lsb_ks_shard_run) over real GPUs:
    torchrun --nproc-per-node G tools/ks_shard_nccl_run.py [logN] [level] [batch] [reps]
Each rank holds only its own limbs; the library owns the NCCL communicator and replays one CUDA graph per call.  Every
rank compares its limbs of ciphertext 0 with the ORACLE's single-device relinearisation (computed on rank 0, broadcast),
times the batched sharded key-switch on the device (max over ranks) and the same driver on a one-rank world (single GPU)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist

import fhe_ref as R
import lattisense_b200 as L
from lattisense_b200.dist import NcclShardedKeySwitch, Ranks, shard_evk, shard_poly


def measure(logN=16, level=24, B=16, reps=20, check=True, rk=None):
    P = {13: R.CKKS_BTP_N13, 15: R.CKKS_BTP_N15, 16: R.CKKS_BTP_N16}[logN]
    rk = rk or Ranks()
    torch.cuda.set_device(rk.local)
    dev = torch.device("cuda", rk.local)
    ctx = L.Context(P.scheme, P.logN, P.q, P.p, P.t, device=rk.local)
    N, l, k = P.N, level + 1, len(P.p)
    beta = -(-l // k)
    rng = np.random.default_rng(5)  # same stream on every rank: identical global inputs
    ct0 = np.stack([np.stack([rng.integers(0, P.q[i], size=N, dtype=np.uint64) for i in range(l)]) for _ in range(3)])
    mods = P.q[:l] + P.p
    evk = np.stack([np.stack([np.stack([rng.integers(0, m, size=N, dtype=np.uint64) for m in mods]) for _ in range(2)]) for _ in range(beta)])
    to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).to(dev)
    sk = NcclShardedKeySwitch(ctx, level, rk.world, rk.rank, max_batch=B, device=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(11)  # items 1..B-1: random residues, identical on every rank (same seed), item 0 = the checked one

    def rand_own(lead, idx):
        t = torch.empty(lead + (max(len(idx), 1), N), dtype=torch.int64, device=dev)
        full = torch.empty(lead + (l, N), dtype=torch.int64, device=dev)
        for i in range(l):
            full[..., i, :] = torch.randint(0, P.q[i], lead + (N,), dtype=torch.int64, device=dev, generator=g)
        if idx:
            t[..., :len(idx), :] = full[..., idx, :]
        return t

    c_own = rand_own((B,), sk.own_q)
    add_own = rand_own((B, 2), sk.own_q)
    if sk.nq_own:
        c_own[0, :sk.nq_own] = to_dev(shard_poly(ct0[2], sk.own_q))
        add_own[0, :, :sk.nq_own] = to_dev(shard_poly(ct0[:2], sk.own_q))
    evk_own = to_dev(shard_evk(evk, level, k, sk.own_q, sk.own_p))
    s = torch.cuda.Stream(device=dev)
    out = torch.zeros((B, 2, max(sk.nq_own, 1), N), dtype=torch.int64, device=dev)
    with torch.cuda.stream(s):
        for _ in range(3):  # eager, capture, replay
            sk.run(c_own, evk_own, add_own, out=out, stream=s.cuda_stream)
        s.synchronize()
    exact = None
    if check:
        want = torch.empty((2, l, N), dtype=torch.int64, device=dev)
        if rk.rank == 0:
            want.copy_(to_dev(R.Oracle(P).relin(ct0, evk)))
        if rk.world > 1:
            dist.broadcast(want, src=0)
        mine = bool(torch.equal(out[0, :, :sk.nq_own], want[:, sk.own_q, :])) if sk.nq_own else True
        exact = rk.sum(1.0 if mine else 0.0) == rk.world
    rk.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        ev0.record(s)
        for _ in range(reps):
            sk.run(c_own, evk_own, add_own, out=out, stream=s.cuda_stream)
        ev1.record(s)
        s.synchronize()
    ms_shard = rk.max(ev0.elapsed_time(ev1) / reps)
    ag_bytes = sk.all_gather_bytes()
    sk.close()
    rk.barrier()
    # single-GPU reference through the same driver (one-rank world) on rank 0
    single = None
    if rk.rank == 0:
        one = NcclShardedKeySwitch(ctx, level, 1, 0, max_batch=B, device=dev)
        c1 = torch.empty((B, l, N), dtype=torch.int64, device=dev)
        a1 = torch.empty((B, 2, l, N), dtype=torch.int64, device=dev)
        for i in range(l):
            c1[:, i, :] = torch.randint(0, P.q[i], (B, N), dtype=torch.int64, device=dev)
            a1[:, :, i, :] = torch.randint(0, P.q[i], (B, 2, N), dtype=torch.int64, device=dev)
        ek = to_dev(evk)
        o1 = torch.empty((B, 2, l, N), dtype=torch.int64, device=dev)
        with torch.cuda.stream(s):
            for _ in range(3):
                one.run(c1, ek, a1, out=o1, stream=s.cuda_stream)
            ev0.record(s)
            for _ in range(reps):
                one.run(c1, ek, a1, out=o1, stream=s.cuda_stream)
            ev1.record(s)
            s.synchronize()
        single = ev0.elapsed_time(ev1) / reps
        one.close()
    rk.barrier()
    res = None
    if rk.rank == 0:
        res = {"workload": "CKKS key-switch n=2^%d level %d (l=%d, k=%d, beta=%d), batch %d, C++ NCCL driver + CUDA graph" % (logN, level, l, k, beta, B),
               "n_gpus": rk.world, "batch": B, "ms_per_batch_sharded_max_over_ranks": ms_shard, "ms_per_batch_single_gpu": single,
               "ms_per_keyswitch_sharded": ms_shard / B, "ms_per_keyswitch_single_gpu": single / B, "speedup_vs_1_gpu": single / ms_shard,
               "ranks_bit_exact_vs_oracle": exact, "all_gather_bytes_per_rank_per_batch": ag_bytes,
               "evk_bytes_per_gpu": int(evk_own.numel() * 8), "evk_bytes_total": int(evk.size * 8)}
    ctx.close()
    return res


if __name__ == "__main__":
    a = [int(x) for x in sys.argv[1:]]
    rk = Ranks()
    r = measure(*(a + [16, 24, 16, 20][len(a):]), check=os.environ.get("KS_CHECK", "1") == "1", rk=rk)
    if r:
        print(json.dumps(r), flush=True)
    rk.close()
