import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import ctypes as C, numpy as np, torch
import fhe_ref as R, lattisense_b200 as L
P = R.CKKS_BTP_N16; dev = torch.device("cuda", 0)
ctx = L.Context(P.scheme, P.logN, P.q, P.p, 0); st = torch.cuda.current_stream().cuda_stream; ctx.stream = st; lib = L.lib()
def limbs(lead, mods):
    t = torch.empty(lead + (len(mods), P.N), dtype=torch.int64, device=dev)
    for i, m in enumerate(mods): t[..., i, :] = torch.randint(0, m, lead + (P.N,), dtype=torch.int64, device=dev)
    return t
for lv in (24, 12):
    l = lv + 1; nrot = 7; B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    ct = limbs((B, 2), P.q[:l]); keys = [limbs((P.beta(lv), 2), P.q[:l] + P.p) for _ in range(nrot)]
    outs = [torch.empty_like(ct) for _ in range(nrot)]
    gals = np.array([R.galois_for_col_step(P.N, i + 1) for i in range(nrot)], dtype=np.uint64)
    pk = (C.c_void_p * nrot)(*[k.data_ptr() for k in keys]); po = (C.c_void_p * nrot)(*[o.data_ptr() for o in outs])
    def plain():
        for i in range(nrot): L._check(lib.lsb_rotate(ctx.h, lv, ct.data_ptr(), keys[i].data_ptr(), lv, int(gals[i]), outs[i].data_ptr(), B, st))
    def hoisted():
        L._check(lib.lsb_rotate_hoisted(ctx.h, lv, ct.data_ptr(), nrot, pk, lv, gals.ctypes.data_as(L.u64p), po, B, st))
    for name, fn in (("plain", plain), ("hoisted", hoisted)):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): fn()
        e1.record(); torch.cuda.synchronize()
        print("level %d B=%d %s: %.3f ms for %d rotations" % (lv, B, name, e0.elapsed_time(e1) / 5, nrot))
    ctx.profile(True); hoisted(); torch.cuda.synchronize()
    print("   hoisted split:", {k: round(ctx.profile_read(i)[0], 3) for i, k in enumerate(["ntt_fwd", "ntt_inv", "basis_ext", "ks_inner", "elementwise"])})
    ctx.profile(False); ctx.profile(True); plain(); torch.cuda.synchronize()
    print("   plain split:  ", {k: round(ctx.profile_read(i)[0], 3) for i, k in enumerate(["ntt_fwd", "ntt_inv", "basis_ext", "ks_inner", "elementwise"])})
    ctx.profile(False)
