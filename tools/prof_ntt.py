"""Small NTT-only driver for ncu captures: python tools/prof_ntt.py [logN] [rows] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import fhe_ref as R
import lattisense_b200 as L

logN = int(sys.argv[1]) if len(sys.argv) > 1 else 14
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 1184
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
P = {14: R.CKKS_N14, 13: R.CKKS_N13, 16: R.CKKS_BTP_N16, 15: R.CKKS_BTP_N15}[logN]
g = L.Context(P.scheme, P.logN, P.q, P.p, 0)
rng = np.random.default_rng(0)
nm = 4
a = np.stack([rng.integers(0, P.q[i % nm], size=P.N, dtype=np.uint64) for i in range(rows)])
d = g.upload(a)
for _ in range(reps):
    g.ntt_dev(d, d, nm, list(range(nm)), False, rows // nm)
    g.ntt_dev(d, d, nm, list(range(nm)), True, rows // nm)
g.sync()
assert np.array_equal(g.download(d, a.shape), a)
print("ok")
