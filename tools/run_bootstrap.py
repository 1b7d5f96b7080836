#!/usr/bin/env python3
"""Run the toy bootstrap pipeline end-to-end on the oracle and print precision.

TEST INFRASTRUCTURE (uses oracle/); not part of the product path.
"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import btp_pipeline as P, fhe_ref as R

if __name__ == "__main__":
    log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    prm = P.toy_params(log_n, 17 if log_n >= 12 else 16)
    orc = R.Oracle(prm); s = P.sparse_secret(orc.N, 32, 5); ev = P.Evaluator(orc, s)
    rng = np.random.default_rng(1); n = orc.N // 2
    z = (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)) * 0.5
    ct0 = P.Ct(orc.encrypt(s, 3, 0, R.ckks_encode(orc, z, 0, 2.0 ** 45)), 2.0 ** 45)
    d = 3 if log_n >= 12 else 2
    t = time.time(); out = P.bootstrap(ev, ct0, cts_depth=d, stc_depth=d); res = ev.decode(out, s)
    print("N=2^%d level %d  max err %.3g  mean precision %.1f bits  %.1f s  %s" % (log_n, out.level, np.abs(res - z).max(), R.mean_precision_bits(res, z), time.time() - t, ev.counts))
