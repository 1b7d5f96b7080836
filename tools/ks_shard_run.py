"""Limb-sharded key-switch over real GPUs: torchrun --nproc-per-node G tools/ks_shard_run.py [logN] [level]
Each rank holds only its own limbs of c2, (c0, c1) and of the relinearisation key; the two all-gathers go over
NCCL/NVLink.  Checks every rank's output limbs against the oracle's single-device relinearisation (small rings), times
the sharded key-switch on the device (max over ranks) and, on rank 0, the single-GPU lsb_relin for comparison."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist
import fhe_ref as R
import lattisense_b200 as L
from lattisense_b200.dist import Ranks, ShardedKeySwitch, shard_evk, shard_poly

logN = int(sys.argv[1]) if len(sys.argv) > 1 else 16
level = int(sys.argv[2]) if len(sys.argv) > 2 else 24
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
check = logN <= 13 or os.environ.get("KS_CHECK") == "1"
P = {13: R.CKKS_BTP_N13, 15: R.CKKS_BTP_N15, 16: R.CKKS_BTP_N16}[logN]
rk = Ranks()
torch.cuda.set_device(rk.local)
ctx = L.Context(P.scheme, P.logN, P.q, P.p, P.t, device=rk.local)
rng = np.random.default_rng(5)              # same stream on every rank: identical global inputs
N, l, k = P.N, level + 1, len(P.p)
beta = -(-l // k)
ct3 = np.stack([np.stack([rng.integers(0, P.q[i], size=N, dtype=np.uint64) for i in range(l)]) for _ in range(3)])
mods = P.q[:l] + P.p
evk = np.stack([np.stack([np.stack([rng.integers(0, m, size=N, dtype=np.uint64) for m in mods]) for _ in range(2)]) for _ in range(beta)])
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()
sk = ShardedKeySwitch(ctx, level, rk.world, rk.rank)
c_own = dev(shard_poly(ct3[2], sk.own_q))
add_own = dev(shard_poly(ct3[:2], sk.own_q))
evk_own = dev(shard_evk(evk, level, k, sk.own_q, sk.own_p))
out = sk.run(c_own, evk_own, add_own)
torch.cuda.synchronize()
ok = None
if check:
    want = R.Oracle(P).relin(ct3, evk)
    ok = bool(np.array_equal(out.cpu().numpy().view(np.uint64), want[:, sk.own_q, :]))
rk.barrier()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(5):
    sk.run(c_own, evk_own, add_own)
rk.barrier()
ev0.record()
for _ in range(reps):
    sk.run(c_own, evk_own, add_own)
ev1.record()
torch.cuda.synchronize()
ms_shard = rk.max(ev0.elapsed_time(ev1) / reps)
# fused compute + collective: INTT kernels store into the peers' buffers over NVLink, two barriers, no NCCL collective
ms_p2p, p2p_same = None, None
if rk.world > 1 and os.environ.get("KS_P2P", "1") == "1":
    sk.enable_p2p()
    out2 = sk.run_p2p(c_own, evk_own, add_own)
    torch.cuda.synchronize()
    p2p_same = bool(torch.equal(out2, out))
    for _ in range(5):
        sk.run_p2p(c_own, evk_own, add_own)
    rk.barrier()
    ev0.record()
    for _ in range(reps):
        sk.run_p2p(c_own, evk_own, add_own)
    ev1.record()
    torch.cuda.synchronize()
    ms_p2p = rk.max(ev0.elapsed_time(ev1) / reps)
    p2p_same = rk.sum(1.0 if p2p_same else 0.0) == rk.world
oks = rk.sum(1.0 if ok in (None, True) else 0.0)
single = None
if rk.rank == 0:
    d3, dk = dev(ct3), dev(evk)
    o2 = torch.empty((2, l, N), dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(5):
        L._check(L.lib().lsb_relin(ctx.h, level, d3.data_ptr(), dk.data_ptr(), level, o2.data_ptr(), 1, st))
    ev0.record()
    for _ in range(reps):
        L._check(L.lib().lsb_relin(ctx.h, level, d3.data_ptr(), dk.data_ptr(), level, o2.data_ptr(), 1, st))
    ev1.record()
    torch.cuda.synchronize()
    single = ev0.elapsed_time(ev1) / reps
    same = bool(torch.equal(o2[:, sk.own_q, :], out))
    print(json.dumps({"workload": "CKKS relinearisation key-switch n=2^%d level %d (l=%d, k=%d, beta=%d)" % (logN, level, l, k, beta),
                      "n_gpus": rk.world, "ms_sharded_max_over_ranks": ms_shard, "ms_single_gpu": single,
                      "ms_sharded_fused_p2p": ms_p2p, "fused_equals_nccl_path_on_all_ranks": p2p_same,
                      "speedup": single / ms_shard, "ranks_bit_exact_vs_oracle": int(oks) if check else None,
                      "rank0_limbs_equal_single_gpu": same,
                      "evk_bytes_per_gpu": int(evk_own.numel() * 8), "evk_bytes_total": int(evk.size * 8),
                      "allgather_bytes_per_rank": int((sk.coeff_own.numel() + sk.accP_own.numel()) * 8)}), flush=True)
rk.barrier()
rk.close()
