"""Bootstrapping (a13) groundwork, CPU only: the special-FFT factorisation the device pipeline will evaluate
(tests/btp_model.py) against the canonical embedding, and its rotation structure against what the reference
frontend computes for the default parameters (tests/golden/btp_rotations.json, made by make_btp_fixture.py from
frontend/bootstrap_params.py)."""
import json
import os

import numpy as np
import pytest

import btp_model as B

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("n", [2, 8, 64, 256])
def test_butterfly_stages_give_the_embedding(n):
    rng = np.random.default_rng(n)
    w = rng.normal(size=n) + 1j * rng.normal(size=n)
    z = B.embedding_matrix(n) @ w
    v = B.bit_reverse(w)
    for lg in range(1, n.bit_length()):
        v = B.apply(B.stage_s2c(n, 1 << lg), v)
    assert np.abs(v - z).max() < 1e-11
    v = z.copy()
    for lg in range(n.bit_length() - 1, 0, -1):
        v = B.apply(B.stage_c2s(n, 1 << lg), v)
    assert np.abs(v - B.bit_reverse(w)).max() < 1e-11


@pytest.mark.parametrize("depth", [1, 2, 3, 4])
def test_merged_matrices_round_trip_and_bsgs(depth):
    n = 128
    rng = np.random.default_rng(depth)
    w = rng.normal(size=n) + 1j * rng.normal(size=n)
    z = B.embedding_matrix(n) @ w
    v = z.copy()
    for m, idx in zip(B.factor_c2s(n, depth), B.dft_index_map(7, depth, True)):
        assert sorted(m) == idx
        direct = B.apply(m, v)
        v = B.apply_bsgs(m, v, B.bsgs_split(idx, n, 2.0))
        assert np.abs(v - direct).max() < 1e-11
    assert np.abs(v - B.bit_reverse(w)).max() < 1e-11  # coefficients, bit-reversed, in the slots
    for m, idx in zip(B.factor_s2c(n, depth), B.dft_index_map(7, depth, False)):
        assert sorted(m) == idx
        v = B.apply_bsgs(m, v, B.bsgs_split(idx, n, 2.0))
    assert np.abs(v - z).max() < 1e-10


def test_rotation_structure_matches_the_reference_frontend():
    with open(os.path.join(HERE, "golden", "btp_rotations.json")) as f:
        ref = json.load(f)
    n = ref["slots"]
    log_slots = n.bit_length() - 1
    union = set()
    for name, c2s in (("cts", True), ("stc", False)):
        mine = B.dft_index_map(log_slots, ref[name]["depth"], c2s)
        assert mine == ref[name]["index"]
        rots = set()
        for idx in mine:
            n1 = B.bsgs_split(idx, n, ref[name]["bsgs_ratio"])
            rots |= {((k // n1) * n1) % n for k in idx} | {k % n1 for k in idx}
        rots.discard(0)
        assert sorted(rots) == ref[name]["rotations"]
        union |= rots
    assert sorted(union) == ref["rotations"] and len(union) == 47  # cfg5: 47 column keys + the row key


def test_eval_mod_model_removes_the_q0_multiples():
    rng = np.random.default_rng(3)
    msg = rng.uniform(-1, 1, size=4096) / 256.0          # message_ratio 256 (custom_task.py:446)
    k = rng.integers(-12, 13, size=4096)                 # q0 multiples left by ModRaise, |I| < K = 16
    got = B.eval_mod_model(msg + k)
    assert np.abs(got - np.sin(2 * np.pi * msg) / (2 * np.pi)).max() < 2.0 ** -22
    assert np.abs(got - msg).max() < 2.0 ** -14          # sin(x) ~ x: the inherent message_ratio error


def _toy_case(log_n=10, levels=16):
    import btp_pipeline as P
    import fhe_ref as R
    prm = P.toy_params(log_n, levels)
    orc = R.Oracle(prm)
    secret = P.sparse_secret(orc.N, 32, 5)
    rng = np.random.default_rng(1)
    z = (rng.uniform(-1, 1, orc.N // 2) + 1j * rng.uniform(-1, 1, orc.N // 2)) * 0.5
    delta = 2.0 ** 45
    ct0 = P.Ct(orc.encrypt(secret, 3, 0, R.ckks_encode(orc, z, 0, delta)), delta)
    return P, R, prm, orc, secret, z, ct0


def test_toy_bootstrap_on_the_oracle_recovers_the_message():
    """N=2^10, 17 primes of 55 bits, sparse secret: a level-0 ciphertext comes back at level 3 with the reference's
    acceptance criterion met (mean real precision >= 10 bits, test_gpu_ckks.cpp:37-43,743-808)."""
    P, R, prm, orc, secret, z, ct0 = _toy_case()
    ev = P.Evaluator(orc, secret)
    out = P.bootstrap(ev, ct0)
    res = ev.decode(out, secret)
    assert out.level == 3
    assert R.mean_precision_bits(res, z) >= 16.0
    assert np.abs(res - z).max() < 1e-4
    assert ev.counts["rotate"] > 150 and ev.counts["mult"] == 28
