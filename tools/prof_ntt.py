"""Small NTT-only driver (ncu captures and A/B timing): python tools/prof_ntt.py [logN] [rows] [reps] [time]
With a 4th argument prints forward / inverse algorithmic GB/s (16*N bytes per limb transform) from wall-clock
around a stream sync over `reps` back-to-back launches."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import fhe_ref as R
import lattisense_b200 as L

logN = int(sys.argv[1]) if len(sys.argv) > 1 else 14
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 1184
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
timing = len(sys.argv) > 4
P = {14: R.CKKS_N14, 13: R.CKKS_N13, 16: R.CKKS_BTP_N16, 15: R.CKKS_BTP_N15, 12: R.CKKS_N12}[logN]
g = L.Context(P.scheme, P.logN, P.q, P.p, 0)
rng = np.random.default_rng(0)
nm = int(os.environ.get("PROF_NMOD", "4"))
mod0 = int(os.environ.get("PROF_MOD0", "0"))
mods = list(range(mod0, mod0 + nm))
allq = P.q + P.p
a = np.stack([rng.integers(0, allq[mods[i % nm]], size=P.N, dtype=np.uint64) for i in range(rows)])
d = g.upload(a)
for _ in range(reps if not timing else 2):
    g.ntt_dev(d, d, nm, mods, False, rows // nm)
    g.ntt_dev(d, d, nm, mods, True, rows // nm)
g.sync()
assert np.array_equal(g.download(d, a.shape), a)
if timing:
    for inv in (False, True):
        g.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            g.ntt_dev(d, d, nm, mods, inv, rows // nm)
        g.sync()
        dt = (time.perf_counter() - t0) / reps
        print("logN=%d rows=%d %s: %.1f us  %.0f GB/s" % (logN, rows, "inv" if inv else "fwd", dt * 1e6, 16.0 * P.N * rows / dt / 1e9))
print("ok")
