"""The bootstrap node's pipeline on the device (reference acceptance unittests/test_gpu_ckks.cpp:743-808).

* limb parity: lsb_bootstrap_run vs the ORACLE run of the same pipeline (tests/btp/btp_oracle.cpp instantiates the product's
  lsb_bootstrap.h with the CPU oracle's ciphertext operations) on the reference's 25 + 5 prime chain at toy ring degrees;
* precision: decrypt + decode of the device result against the encrypted message, >= 10 bits mean real precision (the
  reference's bound) at the toy degree n = 8192 and at the default n = 65536 (dense secret h = 192, sparse h = 32)."""
import os

import numpy as np
import pytest

import btp_ref as B
import fhe_ref as R
import lattisense_b200 as L

pytestmark = pytest.mark.gpu


def _setup(logN, seed, batch=1):
    P = R.Params("btp_n%d" % (1 << logN), R.SCHEME_CKKS, logN, R._BTP_Q, R._BTP_P, scale=2.0 ** 40)
    o = R.Oracle(P)
    spec = B.BtpSpec(P)
    keys = B.BtpKeys(o, spec, seed=seed, h_dense=min(192, P.N // 4))
    rng = np.random.default_rng(seed)
    zs = [rng.uniform(-1, 1, P.N // 2) + 1j * rng.uniform(-1, 1, P.N // 2) for _ in range(batch)]
    cts = np.stack([B.encrypt_message(o, keys, z, 1000 + i) for i, z in enumerate(zs)])
    g = L.Context(P.scheme, P.logN, P.q, P.p, 0)
    bt = L.Bootstrapper(g, list(spec.ints()), list(spec.doubles()))
    bt.set_keys(keys.rlk, keys.gals, keys.glk, keys.swk_dts, keys.swk_std)
    return P, o, spec, keys, zs, cts, g, bt


@pytest.mark.parametrize("logN,batch", [(12, 2), (13, 1)])
def test_bootstrap_limbs_equal_the_oracle_run(logN, batch):
    P, o, spec, keys, zs, cts, g, bt = _setup(logN, 5, batch)
    assert sorted(bt.rotations()) == sorted(B.pipeline_rotations(spec))
    got, st = bt.run(cts)
    want, st_o = B.oracle_bootstrap(o, spec, keys, cts)
    assert got.shape == want.shape == (batch, 2, spec.out_level + 1, P.N)
    assert np.array_equal(got, want), "device bootstrap differs from the oracle run of the same pipeline"
    assert st["rotations"] == st_o["rotations"] and st["mults"] == st_o["mults"]
    for i, z in enumerate(zs):
        bits = R.mean_precision_bits(B.decode(o, keys, got[i], P.scale), z)
        assert bits >= 10.0, bits
    bt.close()
    g.close()


@pytest.mark.parametrize("logN", [13] + ([16] if os.environ.get("LSB_TEST_BTP_N16") else []))
def test_bootstrap_precision(logN):
    """toy_bootstrap (n = 8192) of the reference's GPU tests: level 0 -> level 9, >= 10 bits, and a second run on the cached
    plaintexts is bit-identical.  (n = 65536 runs through the reference boundary in tests/test_runner_gpu.py
    [ckks_btp64k_bootstrap]; LSB_TEST_BTP_N16=1 repeats it here at the kernel-level ABI: 17.6 bits, 117 key switches.)"""
    P, o, spec, keys, zs, cts, g, bt = _setup(logN, 9, 1)
    got, st = bt.run(cts)
    again, _ = bt.run(cts)  # second run on the cached plaintexts: deterministic
    assert np.array_equal(got, again)
    bits = R.mean_precision_bits(B.decode(o, keys, got[0], P.scale), zs[0])
    print("bootstrap n=%d: %.1f bits, %s" % (P.N, bits, st))
    assert got.shape[2] == 10 and bits >= 10.0, bits
    bt.close()
    g.close()
