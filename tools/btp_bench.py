"""CKKS bootstrapping N=2^16 on the device, random residues: time per call and per-kernel-class split.
    python tools/btp_bench.py [batch]      (LSB_BTP_HOIST=0 disables the hoisted baby-step rotations)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import ctypes as C, numpy as np, torch
import fhe_ref as R, lattisense_b200 as L, btp_ref as BT
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
P = getattr(R, os.environ.get("BTP_PARAMS", "CKKS_BTP_N16")); dev = torch.device("cuda", 0)
ctx = L.Context(P.scheme, P.logN, P.q, P.p, 0); st = torch.cuda.current_stream().cuda_stream; ctx.stream = st; lib = L.lib()
def limbs(lead, mods):
    t = torch.empty(lead + (len(mods), P.N), dtype=torch.int64, device=dev)
    for i, m in enumerate(mods): t[..., i, :] = torch.randint(0, m, lead + (P.N,), dtype=torch.int64, device=dev)
    return t
spec = BT.BtpSpec(P); bt = L.Bootstrapper(ctx, list(spec.ints()), list(spec.doubles()))
rots = bt.rotations(); top = len(P.q) - 1
pool = [limbs((P.beta(top), 2), P.q + P.p) for _ in range(4)]
gals = [R.galois_for_col_step(P.N, r) for r in rots] + [R.galois_for_row(P.N)]
garr = np.array(gals, dtype=np.uint64); gptr = (C.c_void_p * len(gals))(*[pool[i % 4].data_ptr() for i in range(len(gals))])
dts = limbs((1, 2), [P.q[0]] + P.p); ct0 = limbs((B, 2), P.q[:1]); out = torch.empty((B, 2, spec.out_level + 1, P.N), dtype=torch.int64, device=dev)
stats = (C.c_longlong * 5)()
run = lambda: L._check(lib.lsb_bootstrap_run(bt.h, ct0.data_ptr(), pool[0].data_ptr(), top, garr.ctypes.data_as(L.u64p), gptr, len(gals), dts.data_ptr(), pool[1].data_ptr(), out.data_ptr(), B, st, stats))
run(); torch.cuda.synchronize()
import time
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(3): run()
e1.record(); t_host = (time.perf_counter() - t0) / 3; torch.cuda.synchronize()
print("bootstrap n=%d batch %d hoist=%s: %.2f ms per call (host enqueue %.2f ms), launches per call %d" % (P.N, B, os.environ.get("LSB_BTP_HOIST", "1"), e0.elapsed_time(e1) / 3, t_host * 1e3, 0))
n0 = ctx.launch_count(); ctx.profile(True); run(); torch.cuda.synchronize()
print("   launches", ctx.launch_count() - n0, {k: round(ctx.profile_read(i)[0], 2) for i, k in enumerate(["ntt_fwd", "ntt_inv", "basis_ext", "ks_inner", "elementwise"])})
