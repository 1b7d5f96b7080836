"""Plain-arithmetic model of the linear part of CKKS bootstrapping — TEST INFRASTRUCTURE.

The reference delegates bootstrapping to Lattigo (`mega_ag_executors_cpu.cpp:451-468` -> CkksBtpContext::bootstrap);
the only in-tree description of the algorithm is the frontend's key-set computation
(`frontend/bootstrap_params.py:96-263`): CoeffsToSlots / SlotsToCoeffs are the special FFT of the canonical
embedding factored into butterfly stages (3 diagonals each: 0, +-2^k), consecutive stages merged into `depth`
matrices, each evaluated baby-step/giant-step.  This module restates that factorisation on complex vectors so
the device pipeline can be checked stage by stage:

  * slot vector z (n = N/2 slots)  <->  packed coefficients w_k = m_k + i*m_{k+n}:  z = U w,  U[j][k] = xi^(5^j * k)
  * SlotsToCoeffs = U   = S_{log n} ... S_1 * BitRev      (decimation in time, strides 1, 2, 4, ...)
  * CoeffsToSlots = U^-1 = BitRev * T_1 ... T_{log n}      (decimation in frequency, strides n/2, n/4, ...)
    with `bit_reversed=False` (custom_task.py:430,455) the two BitRev factors are dropped: they cancel around the
    slot-wise EvalMod.
  * a matrix is a dict {rotation k: diagonal d_k},  (M v)[j] = sum_k d_k[j] * v[(j + k) mod n]

`dft_index_map` re-derives which rotations each merged matrix needs; tests/test_btp_model.py checks it against the
index sets `frontend/bootstrap_params.py` yields for the default parameters (committed under tests/golden/).
"""
import math

import numpy as np


def rot_group(n):
    """5^j mod 4n, j < n (slot j is the evaluation at xi^(5^j), xi = exp(2*pi*i/4n))."""
    g, out = 1, []
    for _ in range(n):
        out.append(g)
        g = g * 5 % (4 * n)
    return np.array(out)


def bit_reverse(v):
    n = len(v)
    bits = n.bit_length() - 1
    idx = np.array([int(format(i, "0%db" % bits)[::-1], 2) if bits else 0 for i in range(n)])
    return np.asarray(v)[idx]


def embedding_matrix(n):
    """U[j][k] = xi^(5^j * k): packed coefficients -> slots."""
    rg = rot_group(n)
    k = np.arange(n)
    return np.exp(2j * np.pi * ((rg[:, None] * k[None, :]) % (4 * n)) / (4 * n))


# ---- diagonal-form matrices -------------------------------------------------------------------
def apply(diags, v):
    n = len(v)
    out = np.zeros(n, dtype=np.complex128)
    for k, d in diags.items():
        out += d * np.roll(v, -k)  # roll(v, -k)[j] = v[j + k]
    return out


def compose(a, b):
    """diagonals of A*B (apply B first):  (AB)_{i+j} += A_i * rot(B_j, i)."""
    n = len(next(iter(a.values())))
    out = {}
    for i, da in a.items():
        for j, db in b.items():
            k = (i + j) % n
            out[k] = out.get(k, 0) + da * np.roll(db, -i)
    return out


def stage_s2c(n, length):
    """One decimation-in-time butterfly stage of U (block size `length`, stride length/2)."""
    h = length // 2
    rg = rot_group(n)
    j = np.arange(n) % length
    tw = np.exp(2j * np.pi * ((rg[j % h] % (4 * length)) * (n // length)) / (4 * n))
    lo = j < h
    d0 = np.where(lo, 1.0, -tw)   # lower half keeps u, upper half takes -w*v[j] ...
    dp = np.where(lo, tw, 0.0)    # ... lower half adds w*v[j+h]
    dm = np.where(lo, 0.0, 1.0)   # ... upper half adds u = v[j-h]
    return {0: d0.astype(np.complex128), h % n: dp.astype(np.complex128), (n - h) % n: dm.astype(np.complex128)} \
        if h % n != (n - h) % n else {0: d0.astype(np.complex128), h % n: (dp + dm).astype(np.complex128)}


def stage_c2s(n, length):
    """Inverse of stage_s2c(n, length) (one decimation-in-frequency stage of U^-1, factor 1/2 included)."""
    h = length // 2
    rg = rot_group(n)
    j = np.arange(n) % length
    twi = np.exp(-2j * np.pi * ((rg[j % h] % (4 * length)) * (n // length)) / (4 * n))
    lo = j < h
    d0 = np.where(lo, 0.5, -0.5 * twi)
    dp = np.where(lo, 0.5, 0.0)
    dm = np.where(lo, 0.0, 0.5 * twi)
    return {0: d0.astype(np.complex128), h % n: dp.astype(np.complex128), (n - h) % n: dm.astype(np.complex128)} \
        if h % n != (n - h) % n else {0: d0.astype(np.complex128), h % n: (dp + dm).astype(np.complex128)}


def merge_plan(log_slots, depth, c2s):
    """How many butterfly stages each of the `depth` matrices absorbs (bootstrap_params.py:104-112)."""
    level, merge = log_slots, [0] * depth
    for i in range(depth):
        d = int(math.ceil(level / (depth - i)))
        merge[i if c2s else depth - i - 1] = d
        level -= d
    return merge


def factor_c2s(n, depth):
    """[M_0, ..., M_{depth-1}], applied in that order to a natural-order slot vector; result = BitRev(U^-1 z)."""
    log_n = n.bit_length() - 1
    mats, level = [], log_n
    for m in merge_plan(log_n, depth, True):
        cur = stage_c2s(n, 1 << level)
        for s in range(1, m):
            cur = compose(stage_c2s(n, 1 << (level - s)), cur)
        mats.append(cur)
        level -= m
    return mats


def factor_s2c(n, depth):
    """[M_0, ...] applied in order to BitRev(w); result = U w."""
    log_n = n.bit_length() - 1
    mats, level = [], log_n
    for m in merge_plan(log_n, depth, False):
        # frontend level `level` <-> rotation 2^(log_n - level): stride 1 first
        cur = stage_s2c(n, 2 << (log_n - level))
        for s in range(1, m):
            cur = compose(stage_s2c(n, 2 << (log_n - level + s)), cur)
        mats.append(cur)
        level -= m
    return mats


def dft_index_map(log_slots, depth, c2s):
    """Rotation set of each merged matrix, derived from the factorisation above (compare
    EncodingMatrixParams.compute_bootstrapping_dft_index_map, bootstrap_params.py:96-139; full-slot case)."""
    n = 1 << log_slots
    out, level = [], log_slots
    for m in merge_plan(log_slots, depth, c2s):
        rots = {0}
        for s in range(m):
            lv = level - s
            r = (1 << (lv - 1)) if c2s else (1 << (log_slots - lv))
            rots = {(a + b) % n for a in rots for b in (0, r, n - r)}
        out.append(sorted(rots))
        level -= m
    return out


# ---- baby-step / giant-step split --------------------------------------------------------------
def bsgs_split(rots, n, ratio):
    """n1 such that #baby / #giant rotations is closest below `ratio` (bootstrap_params.py:188-201)."""
    def counts(n1):
        giant = {((r // n1) * n1) % n for r in rots}
        baby = {r % n1 for r in rots}
        return len(giant) - 1, len(baby) - 1
    n1 = 1
    while n1 < n:
        g, b = counts(n1)
        if g and b / g == ratio:
            return n1
        if g == 0 or b / g > ratio:
            return max(n1 // 2, 1)
        n1 <<= 1
    return 1


def apply_bsgs(diags, v, n1):
    """sum over giant steps g of rot( sum_b rot(d_{g+b}, -g) * rot(v, b), g ): what the device evaluates (the inner
    rotations of v are the hoisted ones, the diagonals are pre-rotated by -g at encode time)."""
    n = len(v)
    groups = {}
    for k, d in diags.items():
        groups.setdefault((k // n1) * n1, []).append((k % n1, d))
    baby = {b: np.roll(v, -b) for g in groups.values() for b, _ in g}
    out = np.zeros(n, dtype=np.complex128)
    for g, items in groups.items():
        acc = np.zeros(n, dtype=np.complex128)
        for b, d in items:
            acc += np.roll(d, g) * baby[b]
        out += np.roll(acc, -g)
    return out


# ---- EvalMod -----------------------------------------------------------------------------------
def chebyshev_coeffs(f, a, b, deg):
    k = np.arange(deg + 1)
    x = np.cos(np.pi * (k + 0.5) / (deg + 1))
    y = f(0.5 * (b - a) * x + 0.5 * (b + a))
    c = np.array([2.0 / (deg + 1) * np.sum(y * np.cos(np.pi * j * (k + 0.5) / (deg + 1))) for j in range(deg + 1)])
    c[0] *= 0.5
    return c


def chebyshev_eval(c, a, b, x):
    t = (2 * x - a - b) / (b - a)
    return np.polynomial.chebyshev.chebval(t, c)


def eval_mod_model(x, K=16, deg=30, double_angle=3):
    """x = (message + q0*I)/q0 with |I| < K  ->  approximately frac part * 2*pi scaled back, i.e. sin(2*pi*x)/(2*pi):
    Cos1 variant (custom_task.py:441-452): cos(2*pi*(x - 1/4)/2^r) by a degree-`deg` Chebyshev interpolant on [-K, K],
    then r double-angle steps cos(2t) = 2cos^2(t) - 1."""
    r = double_angle
    c = chebyshev_coeffs(lambda t: np.cos(2 * np.pi * (t - 0.25) / (1 << r)), -K, K, deg)
    y = chebyshev_eval(c, -K, K, x)
    for _ in range(r):
        y = 2 * y * y - 1
    return y / (2 * np.pi)
