"""CPU checks of the device code's LSB_HD bodies (NTT rounds, basis extension, Barrett/Shoup)
against the oracle — no GPU needed.  These validate indexing/twiddle/range logic of the kernels;
the GPU parity tests proper are in test_gpu_parity.py."""
import ctypes as C

import numpy as np
import pytest

import fhe_ref as R

CASES = [
    (R.CKKS_N12, [0, 3, 4], 256),
    (R.CKKS_N13, [0, 5, 6], 256),
    (R.CKKS_N14, [0, 1, 10, 11], 512),
    (R.BFV_N14, [0, 5, 7], 512),
    (R.BFV_N13, [0, 3], 256),
    (R.CKKS_BTP_N16, [0, 1, 29], 0),
    (R.CKKS_BTP_N15, [0, 12, 25], 0),
]


@pytest.mark.parametrize("params,mids,nt", CASES, ids=[c[0].name for c in CASES])
def test_ntt_rounds_match_oracle(emul, params, mids, nt):
    rng = np.random.default_rng(7)
    o = R.Oracle(params)
    for mid in mids:
        q = o.moduli[mid]
        a = rng.integers(0, q, size=params.N, dtype=np.uint64)
        a[:4] = [0, 1, q - 1, q // 2]
        fwd = a.copy()
        assert emul.emul_ntt(params.logN, q, R.P(fwd), 0, nt) == 0
        assert np.array_equal(fwd, o.ntt(mid, a))
        inv = a.copy()
        assert emul.emul_ntt(params.logN, q, R.P(inv), 1, nt) == 0
        assert np.array_equal(inv, o.ntt(mid, a, inverse=True))
        back = fwd.copy()
        emul.emul_ntt(params.logN, q, R.P(back), 1, nt)
        assert np.array_equal(back, a)


def test_ntt_thread_count_independent(emul):
    rng = np.random.default_rng(3)
    p = R.CKKS_N13
    q = p.q[1]
    a = rng.integers(0, q, size=p.N, dtype=np.uint64)
    outs = []
    for nt in (64, 256, 512, 1024):
        x = a.copy()
        emul.emul_ntt(p.logN, q, R.P(x), 0, nt)
        outs.append(x)
    for x in outs[1:]:
        assert np.array_equal(x, outs[0])


@pytest.mark.parametrize("params", [R.CKKS_N13, R.BFV_N14, R.CKKS_BTP_N13, R.CKKS_N16], ids=lambda p: p.name)
def test_basis_extend_matches_oracle(emul, params):
    """incl. the mixed-pipe CRT (lsb_tables.h bex_choose_mixed): the bootstrap chain's 60/61-bit sources enter the FP64 targets
    as two 32-bit halves, the default n = 65536 set's 46-bit sources whole; worst-case residues ride along"""
    rng = np.random.default_rng(11)
    o = R.Oracle(params)
    mods = np.array(o.moduli, dtype=np.uint64)
    nq, k = len(params.q), params.k
    combos = [(list(range(0, min(k, nq))), [i for i in range(nq) if i >= k] + [nq + j for j in range(k)], 0),
              ([nq + j for j in range(k)], list(range(nq)), 0),
              ([nq - 1], list(range(nq - 1)) + [nq + j for j in range(k)], 1)]
    if params.scheme == R.SCHEME_BFV:
        nqm = o.nmods - nq - k
        combos.append((list(range(nq)), [nq + k + j for j in range(nqm)], 0))
        combos.append(([nq + k + j for j in range(nqm)], list(range(nq)), 0))
    for src, dst, centered in combos:
        src, dst = src[:16], dst[:48]
        x = np.stack([rng.integers(0, o.moduli[i], size=params.N, dtype=np.uint64) for i in src])
        for i, sm in enumerate(src):  # extreme residues in the first columns
            x[i, :6] = [o.moduli[sm] - 1, 0, 1, o.moduli[sm] - 1, o.moduli[sm] // 2, min((1 << 32) - 1, o.moduli[sm] - 1)]
        x[0, 3] = 0
        if not centered:
            want = o.basis_extend(src, x, dst)
        else:
            s = o.moduli[src[0]]
            want = np.zeros((len(dst), params.N), dtype=np.uint64)
            xs = x[0].astype(object)
            for t, d in enumerate(dst):
                pj = o.moduli[d]
                want[t] = np.array([(int(v) % pj) if int(v) < (s >> 1) else (-(s - int(v)) % pj) for v in xs],
                                   dtype=np.uint64)
        sa = (C.c_int * len(src))(*src)
        da = (C.c_int * len(dst))(*dst)
        if not centered and params in (R.CKKS_BTP_N13, R.CKKS_N16):
            mixed = emul.emul_bex_mixed_targets(R.P(mods), len(mods), sa, len(src), da, len(dst))
            assert (mixed & 0xFF) > 0 and not (mixed >> 9), (src, dst, mixed)  # these jobs do use both pipes
            assert bool(mixed & 0x100) == (max(o.moduli[i] for i in src) >= 2 ** 51)
        for force_int in (0, 2):  # default policy (FP64 CRT when every prime is below 2^47), then the integer CRT
            got = np.zeros((len(dst), params.N), dtype=np.uint64)
            assert emul.emul_basis_extend(params.logN, R.P(mods), len(mods), sa, len(src), da, len(dst), centered | force_int,
                                          R.P(x), R.P(got)) == 0
            assert np.array_equal(got, want), (src, dst, centered, force_int)


def test_barrett_and_shoup(emul):
    rng = np.random.default_rng(5)
    for q in [R.CKKS_N14.q[0], R.CKKS_N14.q[1], R.BFV_N14.q[0], R.CKKS_BTP_N16.p[0], 65537]:
        for _ in range(2000):
            a = int(rng.integers(0, q))
            b = int(rng.integers(0, q))
            assert emul.emul_mulmod(a, b, q) == a * b % q
            assert emul.emul_mul_shoup(a, b, q) == a * b % q
            x = int(rng.integers(0, 2 ** 63)) * 2 + int(rng.integers(0, 2))
            assert emul.emul_barrett64(x, q) == x % q
            assert emul.emul_mul_shoup(x, b, q) == x * b % q
        for a, b in [(0, 0), (q - 1, q - 1), (1, q - 1), (q - 1, 1)]:
            assert emul.emul_mulmod(a, b, q) == a * b % q


def test_reduce_wide_up_to_its_bound(emul):
    """shift-Barrett (lsb_math.h reduce_wide): exact for every x < 2^(bitlen(q)-1+64), i.e. sums of n products of
    residues while n*q < 2^63 -- the bound the key-switch inner product and the CRT sums rely on"""
    rng = np.random.default_rng(11)
    for q in [R.CKKS_N14.q[0], R.CKKS_N14.q[1], R.CKKS_N14.p[0], R.BFV_N14.q[0], R.CKKS_BTP_N16.p[0], R.CKKS_BTP_N16.q[0], 65537]:
        bound = 1 << (q.bit_length() - 1 + 64)
        nmax = ((1 << 63) - 1) // q
        xs = [0, q - 1, q, 2 * q, bound - 1, bound - q, min(nmax, 1 << 20) * (q - 1) ** 2]
        xs += [int(rng.integers(0, 2 ** 62)) * int(rng.integers(0, 2 ** 62)) % bound for _ in range(3000)]
        xs += [bound - 1 - int(rng.integers(0, 2 ** 40)) for _ in range(500)]
        for x in xs:
            assert x < bound
            assert emul.emul_reduce_wide(x & (2 ** 64 - 1), x >> 64, q) == x % q, (q, x)


@pytest.mark.parametrize("params", [R.CKKS_N12, R.CKKS_N13, R.CKKS_N14, R.BFV_N13, R.CKKS_BTP_N15, R.CKKS_BTP_N16, R.CKKS_N16],
                         ids=lambda p: p.name)
@pytest.mark.parametrize("kernel", ["emul_ntt_cluster", "emul_ntt_persist"])
def test_cluster_ntt_both_policies_match_oracle(emul, params, kernel):
    """k_ntt_cluster's / k_ntt_persist's round code on the CPU (the persistent kernel reads each block's twiddles from
    its own shared-memory table, lsb_ntt_persist_core.h): the FP64 policy (eligible primes) and the 64-bit policy give
    the oracle's limbs bit for bit, forward and inverse, including extreme residues."""
    run = getattr(emul, kernel)
    o = R.Oracle(params)
    rng = np.random.default_rng(5)
    n_f64 = 0
    mids = range(o.nmods) if params.logN <= 14 else [0, 1, 9, 12, 24, 25, 29]
    for mid in mids:
        q = o.moduli[mid]
        a = rng.integers(0, q, size=params.N, dtype=np.uint64)
        a[:4] = [0, 1, q - 1, q // 2]
        worst = np.full(params.N, q - 1, dtype=np.uint64)
        for f64 in (0, 1, 2) if kernel == "emul_ntt_cluster" else (0, 1):  # 2: FP64 policy with 8-byte twiddles (ArithF64W)
            for vec in (a, worst):
                fwd = vec.copy()
                rc = run(params.logN, q, R.P(fwd), 0, f64)
                if rc == 2:
                    assert f64 >= 1 and q * (params.logN + 1) >= 2 ** 51
                    break
                assert rc == 0
                assert np.array_equal(fwd, o.ntt(mid, vec))
                inv = vec.copy()
                assert run(params.logN, q, R.P(inv), 1, f64) == 0
                assert np.array_equal(inv, o.ntt(mid, vec, inverse=True))
            else:
                n_f64 += 1 if f64 == 1 else 0
    if params.scheme == R.SCHEME_CKKS and params.logN <= 14:
        assert n_f64 == o.nmods  # every default CKKS prime takes the FP64 path
