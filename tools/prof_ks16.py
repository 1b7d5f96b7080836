"""N=2^16 level-24 key-switch (the unit of cfg5) and BFV mult_relin (cfg1): per-kernel-class split by CUDA events, and a
target for ncu captures of the integer-policy kernels:  python tools/prof_ks16.py [ks16|bfv] [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import fhe_ref as R
import lattisense_b200 as L

mode = sys.argv[1] if len(sys.argv) > 1 else "ks16"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
dev = torch.device("cuda", 0)
P = R.CKKS_BTP_N16 if mode == "ks16" else R.BFV_N14
lv = 24 if mode == "ks16" else 2
l = lv + 1
ctx = L.Context(P.scheme, P.logN, P.q, P.p, P.t)
st = torch.cuda.current_stream().cuda_stream
ctx.stream = st
lib = L.lib()


def limbs(lead, mods):
    t = torch.empty(lead + (len(mods), P.N), dtype=torch.int64, device=dev)
    for i, m in enumerate(mods):
        t[..., i, :] = torch.randint(0, m, lead + (P.N,), dtype=torch.int64, device=dev)
    return t


rk = limbs((P.beta(lv), 2), P.q[:l] + P.p)
if mode == "ks16":
    c3 = limbs((B, 3), P.q[:l])
    out = torch.empty((B, 2, l, P.N), dtype=torch.int64, device=dev)
    run = lambda: L._check(lib.lsb_relin(ctx.h, lv, c3.data_ptr(), rk.data_ptr(), lv, out.data_ptr(), B, st))
else:
    a, b = limbs((B, 2), P.q[:l]), limbs((B, 2), P.q[:l])
    out = torch.empty((B, 2, l, P.N), dtype=torch.int64, device=dev)
    run = lambda: L._check(lib.lsb_mult_relin(ctx.h, lv, a.data_ptr(), b.data_ptr(), rk.data_ptr(), lv, out.data_ptr(), B, st))
for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
ctx.profile(True)
run()
torch.cuda.synchronize()
kinds = ["ntt_fwd", "ntt_inv", "basis_ext", "ks_inner", "elementwise"]
prof = {k: ctx.profile_read(i) for i, k in enumerate(kinds)}
ctx.profile(False)
print("%s batch %d: %.3f ms per call = %.4f ms per op" % (mode, B, ms, ms / B))
for k, v in prof.items():
    print("  %-12s %.3f ms  %d launches  %.1f GB/s algorithmic" % (k, v[0], v[1], (v[2] / (v[0] / 1e3) / 1e9) if v[0] > 0 else 0))
