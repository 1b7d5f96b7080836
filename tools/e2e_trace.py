"""Dispatcher timeline (LSB_TRACE=1) of one cfg2 x1024 call through run_fhe_gpu_task on a warm handle (development aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import time, torch
import bench, fhe_ref as R
P = R.CKKS_N14; B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024; l = bench.LEVEL + 1
def limbs(lead, mods):
    t = torch.empty(lead + (len(mods), P.N), dtype=torch.int64).pin_memory()
    for i, m in enumerate(mods): t[..., i, :] = torch.randint(0, m, lead + (P.N,), dtype=torch.int64)
    return t
a, b = limbs((B, 2), P.q[:l]), limbs((B, 2), P.q[:l]); rlk = limbs((P.beta(bench.LEVEL), 2), P.q[:l] + P.p)
out = torch.empty((B, 2, l - 1, P.N), dtype=torch.int64).pin_memory()
arm = bench.RunnerArm(P, B, a, b, rlk, out, 0, "trace")
for _ in range(3): arm.step()
os.environ["LSB_TRACE"] = "1"
t0 = time.perf_counter(); arm.step(); print("call: %.1f ms" % (1e3 * (time.perf_counter() - t0)), arm.host_profile())
os.environ.pop("LSB_TRACE"); arm.close()
