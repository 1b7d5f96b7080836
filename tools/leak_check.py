"""Repeated runs must not grow device or host memory: cfg2 x256 through run_fhe_gpu_task, a bootstrap at n = 8192, kernel-ABI ops
(development aid; prints free device memory and the process RSS before / after)."""
import os, sys, resource
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import ctypes as C, numpy as np, torch
import bench, fhe_ref as R, lattisense_b200 as L, btp_ref as BT
rss = lambda: resource.getrusage(resource.RUSAGE_SELF).ru_maxrss // 1024
free = lambda: torch.cuda.mem_get_info(0)[0] >> 20
P = R.CKKS_N14; B = 256; l = bench.LEVEL + 1
def limbs(lead, mods, n, pin=False, dev=None):
    t = torch.empty(lead + (len(mods), n), dtype=torch.int64, device=dev)
    for i, m in enumerate(mods): t[..., i, :] = torch.randint(0, m, lead + (n,), dtype=torch.int64, device=dev)
    return t.pin_memory() if pin else t
a, b = limbs((B, 2), P.q[:l], P.N, True), limbs((B, 2), P.q[:l], P.N, True); rlk = limbs((P.beta(bench.LEVEL), 2), P.q[:l] + P.p, P.N, True)
out = torch.empty((B, 2, l - 1, P.N), dtype=torch.int64).pin_memory()
arm = bench.RunnerArm(P, B, a, b, rlk, out, 0, "leak")
for _ in range(5): arm.step()
torch.cuda.synchronize(); f0, r0 = free(), rss()
for _ in range(60): arm.step()
torch.cuda.synchronize(); print("run_fhe_gpu_task x60: free device MiB %d -> %d, max RSS MiB %d -> %d" % (f0, free(), r0, rss()))
arm.close()
# create / release cycles
f0, r0 = free(), rss()
for i in range(6):
    arm = bench.RunnerArm(P, 64, a, b, rlk, out, 0, "leak%d" % i); arm.step(); arm.close()
torch.cuda.synchronize(); print("create/run/release x6: free device MiB %d -> %d, max RSS MiB %d -> %d" % (f0, free(), r0, rss()))
# bootstrap at n = 8192
PB = R.CKKS_BTP_N13; dev = torch.device("cuda", 0)
ctx = L.Context(PB.scheme, PB.logN, PB.q, PB.p, 0); st = torch.cuda.current_stream().cuda_stream; ctx.stream = st; lib = L.lib()
spec = BT.BtpSpec(PB); bt = L.Bootstrapper(ctx, list(spec.ints()), list(spec.doubles()))
rots = bt.rotations(); top = len(PB.q) - 1
pool = [limbs((PB.beta(top), 2), PB.q + PB.p, PB.N, dev=dev) for _ in range(2)]
gals = [R.galois_for_col_step(PB.N, r) for r in rots] + [R.galois_for_row(PB.N)]
garr = np.array(gals, dtype=np.uint64); gptr = (C.c_void_p * len(gals))(*[pool[i % 2].data_ptr() for i in range(len(gals))])
dts = limbs((1, 2), [PB.q[0]] + PB.p, PB.N, dev=dev); ct0 = limbs((1, 2), PB.q[:1], PB.N, dev=dev)
ob = torch.empty((1, 2, spec.out_level + 1, PB.N), dtype=torch.int64, device=dev); stats = (C.c_longlong * 5)()
run = lambda: L._check(lib.lsb_bootstrap_run(bt.h, ct0.data_ptr(), pool[0].data_ptr(), top, garr.ctypes.data_as(L.u64p), gptr, len(gals), dts.data_ptr(), pool[1].data_ptr(), ob.data_ptr(), 1, st, stats))
for _ in range(3): run()
torch.cuda.synchronize(); f0, r0 = free(), rss()
for _ in range(40): run()
torch.cuda.synchronize(); print("bootstrap n=8192 x40: free device MiB %d -> %d, max RSS MiB %d -> %d" % (f0, free(), r0, rss()))
