"""Regenerates tests/golden/oracle_digests.json: SHA-256 digests of oracle outputs for fixed seeds.

The reference holds no golden vectors / KATs for this path and its arithmetic library (Lattigo) is
absent, so the vectors are produced by the oracle itself (`python tests/golden/make_golden.py`).
They pin the oracle (and, through the parity tests, the CUDA path) across commits."""
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fhe_ref as R  # noqa: E402


def _h(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()[:32]


def rand_ct(o, rng, level, npoly=2):
    out = np.zeros((npoly, level + 1, o.N), dtype=np.uint64)
    for i in range(level + 1):
        out[:, i, :] = rng.integers(0, o.p.q[i], size=(npoly, o.N), dtype=np.uint64)
    return out


def compute():
    res = {}
    for params in (R.CKKS_N12, R.CKKS_N13, R.BFV_N13):
        o = R.Oracle(params)
        rng = np.random.default_rng(2024)
        s = o.gen_secret(42)
        L = params.max_level
        rlk = o.gen_rlk(s, 43, L)
        gal = R.galois_for_col_step(o.N, 1)
        glk = o.gen_glk(s, 44, L, gal)
        a, b = rand_ct(o, rng, L), rand_ct(o, rng, L)
        d = {"psi": [str(o.psi(i)) for i in range(o.nmods)], "moduli": [str(m) for m in o.moduli]}
        d["ntt"] = _h(o.ntt_limbs(a))
        d["intt"] = _h(o.ntt_limbs(a, inverse=True))
        ct3 = o.mult(a, b)
        d["mult"] = _h(ct3)
        ct2 = o.relin(ct3, rlk)
        d["relin"] = _h(ct2)
        d["rescale"] = _h(o.rescale(ct2))
        d["rotate"] = _h(o.rotate(a, glk, gal))
        d["rlk"] = _h(rlk)
        res[params.name] = d
    return res


if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_digests.json")
    with open(out, "w") as f:
        json.dump(compute(), f, indent=1, sort_keys=True)
    print("wrote", out)
