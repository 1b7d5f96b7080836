"""CKKS bootstrapping composed from the oracle's ciphertext operations at a toy ring degree —
TEST INFRASTRUCTURE and groundwork: it fixes the op sequence, the level / scale bookkeeping and the plaintext
matrices that the device pipeline has to reproduce, and checks them the way the reference checks bootstrapping
(decrypt + decode against the input, unittests/test_gpu_ckks.cpp:743-808).

    ModRaise -> CoeffsToSlots (btp_model.factor_c2s, BSGS) -> split real / imaginary with one conjugation each
    -> EvalMod (Chebyshev interpolant of cos on [-K, K], baby-step / giant-step, double-angle steps)
    -> SlotsToCoeffs (btp_model.factor_s2c)

Every homomorphic step is an oracle op that already has a bit-exact device twin (rotate, conjugate = rotate_row,
ct x pt, ct x ct + relin, rescale, add), so the device version of this file is an orchestration problem, not a new
arithmetic kernel.  Differences from Lattigo's implementation (unverifiable here): multiplication
by i is folded into the neighbouring matrices instead of a monomial product, the polynomial is evaluated with a
depth-6 split instead of the depth-optimal one, and scales are tracked as floats.
"""
import numpy as np
from numpy.polynomial import chebyshev as Cheb

import btp_model as B
import fhe_ref as R


class Ct:
    def __init__(self, data, scale):
        self.data, self.scale = data, float(scale)

    @property
    def level(self):
        return self.data.shape[1] - 1


def _mulmod(x, y, q):
    return ((x.astype(object) * y.astype(object)) % q).astype(np.uint64)


class OracleOps:
    """The ciphertext operations of the pipeline on the CPU oracle.  lattisense_b200.Context offers the same methods
    (rotate, add, sub, mult, relin, rescale, drop_level, mul_plain, add_plain, ntt) on the device, so the pipeline runs
    unchanged on either and the two runs can be compared limb by limb."""

    def __init__(self, orc):
        self.o = orc
        for name in ("rotate", "add", "sub", "mult", "relin", "rescale", "drop_level"):
            setattr(self, name, getattr(orc, name))

    def mul_plain(self, ct, pt):
        out = np.zeros_like(ct)
        for p in range(ct.shape[0]):
            for i in range(ct.shape[1]):
                out[p, i] = _mulmod(ct[p, i], pt[i], self.o.p.q[i])
        return out

    def add_plain(self, ct, pt):
        out = ct.copy()
        for i in range(ct.shape[1]):
            q = self.o.p.q[i]
            out[0, i] = ((ct[0, i].astype(object) + pt[i].astype(object)) % q).astype(np.uint64)
        return out

    def ntt(self, a, mod_idx, inverse=False):
        return np.stack([self.o.ntt_limbs(a[p], inverse=inverse) for p in range(a.shape[0])])


class Evaluator:
    """Scale-tracking wrapper over a set of ciphertext operations (`ops`: OracleOps or the device Context).  Holds the
    secret only to make keys on demand and to decode (test code); keys, encoding and decoding always use the oracle."""

    def __init__(self, orc, secret, seed=11, ops=None):
        self.o, self.s, self.seed = orc, secret, seed
        self.ops = ops if ops is not None else OracleOps(orc)
        self.N, self.n = orc.N, orc.N // 2
        self.q = orc.p.q
        self.top = len(self.q) - 1
        self.rlk = orc.gen_rlk(secret, seed, self.top)
        self.glk = {}
        self.counts = {"rotate": 0, "mult": 0, "mul_plain": 0, "rescale": 0}

    def _key(self, gal):
        if gal not in self.glk:
            # the key's randomness depends on the Galois element only, so that two runs that ask for keys in a
            # different order (oracle ops one by one, device op per matrix) use identical keys
            self.glk[gal] = self.o.gen_glk(self.s, self.seed + 1 + gal % 1000003, self.top, gal)
        return self.glk[gal]

    def rotate(self, ct, k):
        k %= self.n
        if k == 0:
            return ct
        gal = R.galois_for_col_step(self.N, k)
        self.counts["rotate"] += 1
        return Ct(self.ops.rotate(ct.data, self._key(gal), gal), ct.scale)

    def conj(self, ct):
        gal = R.galois_for_row(self.N)
        self.counts["rotate"] += 1
        return Ct(self.ops.rotate(ct.data, self._key(gal), gal), ct.scale)

    def add(self, a, b):
        assert a.level == b.level and abs(a.scale / b.scale - 1) < 1e-7, (a.level, b.level, a.scale, b.scale)
        return Ct(self.ops.add(a.data, b.data), a.scale)

    def sub(self, a, b):
        assert a.level == b.level and abs(a.scale / b.scale - 1) < 1e-7
        return Ct(self.ops.sub(a.data, b.data), a.scale)

    def drop_to(self, ct, level):
        d = ct.data
        while d.shape[1] - 1 > level:
            d = self.ops.drop_level(d)
        return Ct(d, ct.scale)

    def rescale(self, ct):
        self.counts["rescale"] += 1
        return Ct(self.ops.rescale(ct.data), ct.scale / self.q[ct.level])

    def mul(self, a, b):
        lv = min(a.level, b.level)
        a, b = self.drop_to(a, lv), self.drop_to(b, lv)
        self.counts["mult"] += 1
        return Ct(self.ops.relin(self.ops.mult(a.data, b.data), self.rlk), a.scale * b.scale)

    def _const_pt(self, k, level):
        return np.stack([np.full(self.N, k % self.q[i], dtype=np.uint64) for i in range(level + 1)])

    def mul_vec(self, ct, vec, pt_scale):
        """ct x encode(vec) (n complex slot values) at plaintext scale pt_scale."""
        pt = R.ckks_encode(self.o, vec, ct.level, pt_scale)
        self.counts["mul_plain"] += 1
        return Ct(self.ops.mul_plain(ct.data, pt), ct.scale * pt_scale)

    def mul_const(self, ct, c, const_scale):
        # a constant polynomial is the same constant in every NTT slot
        return Ct(self.ops.mul_plain(ct.data, self._const_pt(int(round(c * const_scale)), ct.level)), ct.scale * const_scale)

    def mul_const_int(self, ct, k):
        return Ct(self.ops.mul_plain(ct.data, self._const_pt(k, ct.level)), ct.scale)

    def add_const(self, ct, c):
        return Ct(self.ops.add_plain(ct.data, self._const_pt(int(round(c * ct.scale)), ct.level)), ct.scale)

    def decode(self, ct, secret):
        pt = self.o.ntt_limbs(self.o.decrypt(secret, ct.data), inverse=True)
        use = min(ct.level + 1, 3)
        coeffs, _ = R.crt_centered(pt[:use], self.q[:use])
        return R.ckks_decode_coeffs(coeffs, self.N, ct.scale)

    # ---- linear transform, baby-step / giant-step ----
    def linear_plan(self, level, diags, n1, pt_scale, factor=1.0):
        """[(k, encoded diagonal k rotated by -(giant step of k))] -- what a bootstrapping context precomputes."""
        return [(k, R.ckks_encode(self.o, np.roll(d, (k // n1) * n1) * factor, level, pt_scale)) for k, d in sorted(diags.items())]

    def linear(self, ct, diags, n1, pt_scale, factor=1.0):
        """sum_g rot( sum_b pt_{g+b} * rot(ct, b), g ), then one rescale."""
        plan = self.linear_plan(ct.level, diags, n1, pt_scale, factor)
        if hasattr(self.ops, "linear_transform"):  # one device operation (lsb_linear_transform)
            rots = {r for k, _ in plan for r in (k % n1, (k // n1) * n1) if r}
            keys = {r: self._key(R.galois_for_col_step(self.N, r)) for r in rots}
            self.counts["rotate"] += len(rots)
            self.counts["mul_plain"] += len(plan)
            self.counts["rescale"] += 1
            return Ct(self.ops.linear_transform(ct.data, plan, n1, keys), ct.scale * pt_scale / self.q[ct.level])
        groups = {}
        for k, pt in plan:
            groups.setdefault((k // n1) * n1, []).append((k % n1, pt))
        baby = {b: self.rotate(ct, b) for items in groups.values() for b, _ in items}
        total = None
        for g, items in sorted(groups.items()):
            acc = None
            for b, pt in items:
                self.counts["mul_plain"] += 1
                term = Ct(self.ops.mul_plain(baby[b].data, pt), ct.scale * pt_scale)
                acc = term if acc is None else self.add(acc, term)
            acc = self.rotate(acc, g)
            total = acc if total is None else self.add(total, acc)
        return self.rescale(total)


def mod_raise(ev, ct0):
    """level-0 ciphertext -> top level: the centred residues mod q0 re-read modulo every prime (Lattigo `ModRaise`).
    The result encrypts Delta*m + q0*I with a small integer polynomial I (|I| grows with the secret's weight)."""
    if hasattr(ev.ops, "mod_raise"):  # the device has it as one operation (lsb_mod_raise)
        return Ct(ev.ops.mod_raise(ct0.data, ev.top), ct0.scale)
    q0 = ev.q[0]
    coef = ev.ops.ntt(ct0.data, [0], inverse=True)
    out = np.zeros((2, ev.top + 1, ev.N), dtype=np.uint64)
    for p in range(2):
        c = coef[p, 0].astype(object)
        c = np.where(c > q0 // 2, c - q0, c)
        for i, qi in enumerate(ev.q):
            out[p, i] = (c % qi).astype(np.uint64)
    return Ct(ev.ops.ntt(out, list(range(ev.top + 1))), ct0.scale)


def cheb_eval(ev, t, coeffs):
    """p(t) = sum c_j T_j(t), deg <= 31, |t| <= 1: T_1..T_8, T_16 by the product formulas, then
    p = (A1*T_8 + A0)*T_16 + (B1*T_8 + B0) with the four degree-<8 parts as constant combinations."""
    S = t.scale
    T = {1: t}

    def dbl(a):  # T_2k = 2 T_k^2 - 1
        return ev.add_const(ev.rescale(ev.mul_const_int(ev.mul(a, a), 2)), -1.0)

    def prod(a, b, c):  # T_{i+j} = 2 T_i T_j - T_{|i-j|}
        m = ev.rescale(ev.mul_const_int(ev.mul(a, b), 2))
        return ev.sub(m, _rescale_to(ev, c, m))

    T[2] = dbl(T[1])
    T[3] = prod(T[1], T[2], T[1])
    T[4] = dbl(T[2])
    T[5] = prod(T[2], T[3], T[1])
    T[6] = dbl(T[3])
    T[7] = prod(T[3], T[4], T[1])
    T[8] = dbl(T[4])
    T[16] = dbl(T[8])
    c = np.zeros(32)
    c[:len(coeffs)] = coeffs
    A, Bp = Cheb.chebdiv(c, [0] * 16 + [1])
    parts = []
    for poly in (A, Bp):
        hi, lo = Cheb.chebdiv(np.pad(poly, (0, max(0, 16 - len(poly)))), [0] * 8 + [1])
        parts.append((hi, lo))
    base = T[7].level

    def small(cs):  # sum_{j<8} cs[j] T_j  -> level base-1, scale S
        cs = np.pad(cs, (0, max(0, 8 - len(cs))))[:8]
        acc = None
        for j in range(1, 8):
            if abs(cs[j]) < 1e-300:
                continue
            x = ev.drop_to(T[j], base)
            term = ev.mul_const(x, cs[j], ev.q[base] * S / x.scale)
            acc = term if acc is None else ev.add(acc, term)
        acc = ev.rescale(acc)
        return ev.add_const(acc, cs[0])

    def times(a, tk):
        m = ev.rescale(ev.mul(a, tk))
        return m

    halves = []
    for hi, lo in parts:
        h = small(hi)
        l = small(lo)
        m = times(h, T[8])
        halves.append(ev.add(m, _rescale_to(ev, l, m)))
    top = times(halves[0], T[16])
    return ev.add(top, _rescale_to(ev, halves[1], top))


def _rescale_to(ev, x, like):
    """bring x to like's level; scales are equal up to the primes' distance from each other (asserted in add)."""
    x = ev.drop_to(x, like.level)
    return Ct(x.data, like.scale) if abs(x.scale / like.scale - 1) < 1e-6 else x




# ---- parameters and the pipeline ----------------------------------------------------------------
def _is_prime(n):
    if n < 2:
        return False
    for p in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
        if n % p == 0:
            return n == p
    d, r = n - 1, 0
    while d % 2 == 0:
        d //= 2
        r += 1
    for a in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(r - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def ntt_primes(bits, two_n, count, skip=()):
    """`count` primes = 1 mod 2N alternating around 2^bits (closest first), so that they all lie within ~2^-35 of it."""
    out, k = [], 0
    base = (1 << bits) + 1
    while len(out) < count:
        k += 1
        for cand in (base + k * two_n, base - k * two_n):
            if len(out) < count and cand not in skip and _is_prime(cand):
                out.append(cand)
    return out


def toy_params(log_n=10, levels=16, bits=55):
    q = ntt_primes(bits, 2 << log_n, levels + 1)
    p = ntt_primes(bits + 1, 2 << log_n, 3)
    return R.Params("BTP_TOY_N%d" % log_n, R.SCHEME_CKKS, log_n, q, p, 0, float(1 << 45))


def sparse_secret(N, weight, seed):
    rng = np.random.default_rng(seed)
    s = np.zeros(N, dtype=np.int8)
    idx = rng.choice(N, size=weight, replace=False)
    s[idx] = rng.choice(np.array([-1, 1], dtype=np.int8), size=weight)
    return s


def coeffs_to_slots(ev, ct, depth, K):
    """-> (t_re, t_im): slots hold coefficient k / (q0*K) resp. coefficient k+n / (q0*K), bit-reversed order."""
    n = ev.n
    mats = B.factor_c2s(n, depth)
    idx = B.dft_index_map(n.bit_length() - 1, depth, True)
    v = ct
    for m, ix in zip(mats[:-1], idx[:-1]):
        v = ev.linear(v, m, B.bsgs_split(ix, n, 2.0), ev.q[v.level])
    n1 = B.bsgs_split(idx[-1], n, 2.0)
    a = ev.linear(v, mats[-1], n1, ev.q[v.level], 0.5 / K)
    b = ev.linear(v, mats[-1], n1, ev.q[v.level], -0.5j / K)
    return ev.add(a, ev.conj(a)), ev.add(b, ev.conj(b))


def eval_mod(ev, t, K, deg, double_angle):
    """t = x/K in [-1, 1]  ->  sin(2*pi*x)  (Cos1 with `double_angle` doublings, custom_task.py:441-452)."""
    r = double_angle
    c = B.chebyshev_coeffs(lambda u: np.cos(2 * np.pi * (K * u - 0.25) / (1 << r)), -1.0, 1.0, deg)
    y = cheb_eval(ev, t, c)
    for _ in range(r):
        y = ev.add_const(ev.rescale(ev.mul_const_int(ev.mul(y, y), 2)), -1.0)
    return y


def slots_to_coeffs(ev, y_re, y_im, depth, factor):
    n = ev.n
    mats = B.factor_s2c(n, depth)
    idx = B.dft_index_map(n.bit_length() - 1, depth, False)
    n1 = B.bsgs_split(idx[0], n, 2.0)
    v = ev.add(ev.linear(y_re, mats[0], n1, ev.q[y_re.level], factor), ev.linear(y_im, mats[0], n1, ev.q[y_im.level], 1j * factor))
    for m, ix in zip(mats[1:], idx[1:]):
        v = ev.linear(v, m, B.bsgs_split(ix, n, 2.0), ev.q[v.level])
    return v


def bootstrap(ev, ct0, K=16, deg=30, double_angle=3, cts_depth=2, stc_depth=2):
    """level-0 ciphertext of scale Delta -> ciphertext of the same message at level top - depth, scale Delta."""
    delta, q0 = ct0.scale, float(ev.q[0])
    raised = mod_raise(ev, ct0)
    raised.scale = q0                      # read the plaintext Delta*m + q0*I in units of q0:  x = m*Delta/q0 + I
    t_re, t_im = coeffs_to_slots(ev, raised, cts_depth, K)
    y_re = eval_mod(ev, t_re, K, deg, double_angle)
    y_im = eval_mod(ev, t_im, K, deg, double_angle)
    # sin(2 pi x) ~ 2 pi m Delta / q0: the factor back to m goes into the first SlotsToCoeffs matrix
    out = slots_to_coeffs(ev, y_re, y_im, stc_depth, q0 / (2 * np.pi * delta))
    return out
