"""mult_relin_rescale and rotate at the top level of the reference frontend's default CKKS n = 65536 set (development aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import fhe_ref as R, lattisense_b200 as L
P = R.CKKS_N16; B = 16; dev = torch.device("cuda", 0)
c = L.Context(P.scheme, P.logN, P.q, P.p, 0); st = torch.cuda.current_stream().cuda_stream; c.stream = st; lib = L.lib()
def limbs(lead, mods):
    t = torch.empty(lead + (len(mods), P.N), dtype=torch.int64, device=dev)
    for i, m in enumerate(mods): t[..., i, :] = torch.randint(0, m, lead + (P.N,), dtype=torch.int64, device=dev)
    return t
l = len(P.q); xa, xb = limbs((B, 2), P.q), limbs((B, 2), P.q); k = limbs((P.beta(l - 1), 2), P.q + P.p)
xo = torch.empty((B, 2, l - 1, P.N), dtype=torch.int64, device=dev)
run = lambda: L._check(lib.lsb_mult_relin_rescale(c.h, l - 1, xa.data_ptr(), xb.data_ptr(), k.data_ptr(), l - 1, xo.data_ptr(), B, st))
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): run()
e1.record(); torch.cuda.synchronize()
print("default n=65536 level %d mult_relin_rescale batch %d: %.3f ms per op" % (l - 1, B, e0.elapsed_time(e1) / 5 / B))
c.profile(True); run(); torch.cuda.synchronize()
print({kk: round(c.profile_read(i)[0], 2) for i, kk in enumerate(["ntt_fwd", "ntt_inv", "basis_ext", "ks_inner", "elementwise"])})
