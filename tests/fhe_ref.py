"""Test support: ctypes binding of the CPU oracle (oracle/_build/liborc.so), the reference's parameter
sets, and client-side encode/decode written with numpy + Python integers.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module; the product
package (lattisense_b200/) never does.

Parameter sets: /root/reference/frontend/parameter.json:26-40 (BFV n=16384), :95-114 (CKKS n=16384),
unittests/fixture.hpp:74-103 (custom n=8192 sets), frontend/custom_task.py:302-335,387-420 (BTP chain).
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_DIR = os.path.join(ROOT, "oracle")
ORC_SO = os.path.join(ORC_DIR, "_build", "liborc.so")
ORC_FAST_SO = os.path.join(ORC_DIR, "_build", "liborc_fast.so")

SCHEME_BFV, SCHEME_CKKS = 0, 1

u64p = C.POINTER(C.c_uint64)
i8p = C.POINTER(C.c_int8)
i32p = C.POINTER(C.c_int)


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORC_DIR])
    return ORC_SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORC_SO):
            build_oracle()
        _lib = C.CDLL(ORC_SO)
        _lib.orc_create.restype = C.c_void_p
        _lib.orc_create.argtypes = [C.c_int, C.c_int, u64p, C.c_int, u64p, C.c_int, C.c_uint64]
        _lib.orc_destroy.argtypes = [C.c_void_p]
        _lib.orc_modulus.restype = C.c_uint64
        _lib.orc_modulus.argtypes = [C.c_void_p, C.c_int]
        _lib.orc_psi.restype = C.c_uint64
        _lib.orc_psi.argtypes = [C.c_void_p, C.c_int]
        for name in ("orc_num_moduli", "orc_num_qmul"):
            getattr(_lib, name).argtypes = [C.c_void_p]
        _lib.orc_level_qmul.argtypes = [C.c_void_p, C.c_int]
    return _lib


def P(a):
    """numpy uint64 array -> pointer"""
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


# ---------------------------------------------------------------------------------------------
# Parameter sets
# ---------------------------------------------------------------------------------------------
class Params:
    def __init__(self, name, scheme, logN, q, p, t=0, scale=0.0):
        self.name, self.scheme, self.logN, self.N = name, scheme, logN, 1 << logN
        self.q, self.p, self.t, self.scale = list(q), list(p), t, float(scale)
        self.max_level = len(q) - 1

    @property
    def k(self):
        return len(self.p)

    def beta(self, level):
        return (level + 1 + self.k - 1) // self.k


BFV_N14 = Params("bfv_n16384", SCHEME_BFV, 14,
                 [72057594038321153, 36028797019389953, 36028797019488257, 18014398508400641, 18014398510661633,
                  18014398508138497],
                 [36028797020209153, 36028797017456641], t=65537)
BFV_N13 = Params("bfv_custom_n8192", SCHEME_BFV, 13, [0x3FFFFFFFEF8001, 0x4000000011C001, 0x40000000120001],
                 [0x7FFFFFFFFB4001], t=65537)
CKKS_N14 = Params("ckks_n16384", SCHEME_CKKS, 14,
                  [35184372121601, 17179967489, 17179672577, 17180262401, 17180295169, 17179410433, 17180393473,
                   17180557313, 17180950529, 17178525697],
                  [8796092858369, 8796092792833], scale=2.0 ** 34)
CKKS_N13 = Params("ckks_custom_n8192", SCHEME_CKKS, 13,
                  [0x1FFFEC001, 0x3FFF4001, 0x3FFE8001, 0x40020001, 0x40038001, 0x3FFC0001], [0x800004001],
                  scale=2.0 ** 30)
_BTP_Q = [0x10000000006E0001, 0x10000140001, 0xFFFFE80001, 0xFFFFC40001, 0x100003E0001, 0xFFFFB20001, 0x10000500001,
          0xFFFF940001, 0xFFFF8A0001, 0xFFFF820001, 0x7FFFE60001, 0x7FFFE40001, 0x7FFFE00001, 0xFFFFFFFFF840001,
          0x1000000000860001, 0xFFFFFFFFF6A0001, 0x1000000000980001, 0xFFFFFFFFF5A0001, 0x1000000000B00001,
          0x1000000000CE0001, 0xFFFFFFFFF2A0001, 0x100000000060001, 0xFFFFFFFFF00001, 0xFFFFFFFFD80001,
          0x1000000002A0001]
_BTP_P = [0x1FFFFFFFFFE00001, 0x1FFFFFFFFFC80001, 0x1FFFFFFFFFB40001, 0x1FFFFFFFFF500001, 0x1FFFFFFFFF420001]
CKKS_BTP_N16 = Params("ckks_btp_n65536", SCHEME_CKKS, 16, _BTP_Q, _BTP_P, scale=2.0 ** 40)
CKKS_BTP_N13 = Params("ckks_btp_toy_n8192", SCHEME_CKKS, 13, _BTP_Q, _BTP_P, scale=2.0 ** 40)
CKKS_BTP_N15 = Params("ckks_btp_n32768", SCHEME_CKKS, 15, _BTP_Q, _BTP_P, scale=2.0 ** 40)
# the reference frontend's default CKKS sets for n = 32768 and n = 65536 (frontend/parameter.json "CKKS"/"32768", "65536";
# the file says max_level 33 for the latter but lists 30 Q primes -- SURVEY.md fact 8 -- so the chain here has levels 0..29).
# At its top levels the n = 65536 set has beta = 8 digits over 34 limbs = 272 extended limbs: more than one launch group
# of the key-switch ModUp (Context::ks_digit_group).
CKKS_N15 = Params("ckks_n32768", SCHEME_CKKS, 15,
                  [0x4000000120001, 0x10000140001, 0xFFFFE80001, 0x10000290001, 0xFFFFC40001, 0x100003E0001, 0x10000470001,
                   0x100004B0001, 0xFFFFB20001, 0x10000500001, 0x10000650001, 0xFFFF940001, 0xFFFF8A0001, 0xFFFF820001,
                   0xFFFF780001, 0x10000890001, 0xFFFF750001, 0x10000960001],
                  [0x40000001B0001, 0x3FFFFFFDF0001, 0x4000000270001], scale=2.0 ** 40)
CKKS_N16 = Params("ckks_n65536", SCHEME_CKKS, 16,
                  [0x80000000080001, 0x2000000A0001, 0x2000000E0001, 0x1FFFFFC20001, 0x200000440001, 0x200000500001,
                   0x200000620001, 0x1FFFFF980001, 0x2000006A0001, 0x1FFFFF7E0001, 0x200000860001, 0x200000A60001,
                   0x200000AA0001, 0x200000B20001, 0x200000FE0001, 0x1FFFFEDE0001, 0x1FFFFECA0001, 0x1FFFFEB40001,
                   0x200001520001, 0x1FFFFE760001, 0x2000019A0001, 0x1FFFFE640001, 0x200001A00001, 0x1FFFFE520001,
                   0x200001E80001, 0x1FFFFE0C0001, 0x1FFFFDEE0001, 0x200002480001, 0x1FFFFDB60001, 0x200002560001],
                  [0x80000000440001, 0x7FFFFFFFBA0001, 0x80000000500001, 0x7FFFFFFFAA0001], scale=2.0 ** 45)
# the remaining default sets of frontend/parameter.json (BFV 8192 / 16384 and CKKS 8192 / 16384 are BFV_N13, BFV_N14,
# CKKS_N13, CKKS_N14 above)
BFV_N12 = Params("bfv_n4096", SCHEME_BFV, 12, [0x7FFFFEC001, 0x8000016001], [0x40002001], t=65537)
CKKS_N12_DEFAULT = Params("ckks_n4096", SCHEME_CKKS, 12, [0x200000E001, 0x100006001], [0x3FFFFEA001], scale=2.0 ** 32)
BFV_N15 = Params("bfv_n32768", SCHEME_BFV, 15,
                 [0x7FFFFFFFFE70001, 0x7FFFFFFFFE10001, 0x7FFFFFFFFCC0001, 0x400000000270001, 0x400000000350001,
                  0x400000000360001, 0x3FFFFFFFFC10001, 0x3FFFFFFFFBE0001, 0x3FFFFFFFFBD0001, 0x4000000004D0001,
                  0x400000000570001, 0x400000000660001],
                 [0xFFFFFFFFFFC0001, 0x10000000001D0001, 0x10000000006E0001], t=65537)
# parity-only: 25 Q limbs with 2-limb digits (beta = 13, 351 extended limbs at the top level) on a small ring -- the
# grouped ModUp with groups of unequal size and a short last digit
CKKS_MANYDIGIT_N13 = Params("ckks_many_digits_n8192", SCHEME_CKKS, 13, _BTP_Q, _BTP_P[:2], scale=2.0 ** 40)
# small-ring variants (not NTT-friendly for every reference prime, so use the n=8192 sets' primes: = 1 mod 2^14)
CKKS_N12 = Params("ckks_small_n4096", SCHEME_CKKS, 12, CKKS_N13.q[:4], CKKS_N13.p, scale=2.0 ** 30)

# parity-only set with 3-limb key-switch digits (alpha = 3; digits of 3, 2 and 1 limbs depending on the level):
# exercises the generic-beta / generic-alpha code paths with primes below 2^47 (moduli borrowed from CKKS_N14)
CKKS_WIDE_N13 = Params("ckks_wide_digit_n8192", SCHEME_CKKS, 13, CKKS_N14.q[:7], CKKS_N14.q[7:10], scale=2.0 ** 34)

ALL_PARAMS = {p.name: p for p in (BFV_N14, BFV_N13, CKKS_N14, CKKS_N13, CKKS_BTP_N16, CKKS_BTP_N13, CKKS_BTP_N15, CKKS_N12,
                                     CKKS_WIDE_N13)}


# ---------------------------------------------------------------------------------------------
# Optimised CPU baseline (oracle/orc_fast.hpp): same algorithms with Shoup/Barrett lazy arithmetic, timed by bench.py's
# reference arm; tests assert its limbs equal the simple oracle's
# ---------------------------------------------------------------------------------------------
_flib = None


def fast_lib():
    global _flib
    if _flib is None:
        if not os.path.exists(ORC_FAST_SO):
            build_oracle()
        _flib = C.CDLL(ORC_FAST_SO)
        _flib.orcf_create.restype = C.c_void_p
        _flib.orcf_create.argtypes = [C.c_int, C.c_int, u64p, C.c_int, u64p, C.c_int, C.c_uint64]
        _flib.orcf_destroy.argtypes = [C.c_void_p]
        _flib.orcf_time_mult_relin_rescale.restype = C.c_double
    return _flib


class FastOracle:
    """CKKS mult / relin / rescale / rotate of the optimised CPU baseline (same array conventions as Oracle)"""

    def __init__(self, params):
        self.p, self.N = params, params.N
        q = np.array(params.q, dtype=np.uint64)
        pp = np.array(params.p, dtype=np.uint64)
        self.h = fast_lib().orcf_create(params.scheme, params.logN, P(q), len(q), P(pp), len(pp), params.t)
        if not self.h:
            raise RuntimeError("orcf_create failed for " + params.name)

    def __del__(self):
        if getattr(self, "h", None):
            fast_lib().orcf_destroy(C.c_void_p(self.h))
            self.h = None

    @property
    def hp(self):
        return C.c_void_p(self.h)

    def ntt(self, midx, a, inverse=False):
        out = np.ascontiguousarray(a, dtype=np.uint64).copy()
        fast_lib().orcf_ntt(self.hp, midx, P(out), 1 if inverse else 0)
        return out

    def mult_relin(self, a, b, rlk, rescale=False):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = np.ascontiguousarray(b, dtype=np.uint64)
        l = a.shape[1]
        key_level = rlk.shape[2] - self.p.k - 1
        o = np.zeros((2, l - 1 if rescale else l, self.N), dtype=np.uint64)
        fn = fast_lib().orcf_ckks_mult_relin_rescale if rescale else fast_lib().orcf_ckks_mult_relin
        fn(self.hp, l - 1, P(a), P(b), P(rlk), key_level, P(o))
        return o

    def time_mult_relin_rescale(self, a, b, rlk, reps):
        l = a.shape[1]
        key_level = rlk.shape[2] - self.p.k - 1
        o = np.zeros((2, l - 1, self.N), dtype=np.uint64)
        return fast_lib().orcf_time_mult_relin_rescale(self.hp, l - 1, P(a), P(b), P(rlk), key_level, P(o), reps)

    def rotate(self, ct, glk, gal):
        ct = np.ascontiguousarray(ct, dtype=np.uint64)
        l = ct.shape[1]
        key_level = glk.shape[2] - self.p.k - 1
        o = np.zeros((2, l, self.N), dtype=np.uint64)
        fast_lib().orcf_ckks_rotate(self.hp, l - 1, P(ct), P(glk), key_level, C.c_uint64(gal), P(o))
        return o


# ---------------------------------------------------------------------------------------------
# Oracle wrapper
# ---------------------------------------------------------------------------------------------
class Oracle:
    def __init__(self, params):
        self.p = params
        L = lib()
        q = np.array(params.q, dtype=np.uint64)
        pp = np.array(params.p, dtype=np.uint64)
        self.h = L.orc_create(params.scheme, params.logN, P(q), len(q), P(pp), len(pp), params.t)
        if not self.h:
            raise RuntimeError("orc_create failed for " + params.name)
        self.N = params.N
        self.nmods = L.orc_num_moduli(self.h)
        self.moduli = [L.orc_modulus(self.h, i) for i in range(self.nmods)]

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(C.c_void_p(self.h))
            self.h = None

    @property
    def hp(self):
        return C.c_void_p(self.h)

    def psi(self, midx):
        return lib().orc_psi(self.hp, midx)

    def ntt(self, midx, a, inverse=False):
        out = np.ascontiguousarray(a, dtype=np.uint64).copy()
        lib().orc_ntt(self.hp, midx, P(out), 1 if inverse else 0)
        return out

    def ntt_limbs(self, a, inverse=False, midx0=0):
        """a: [..., l, N] -> per-limb transform with moduli midx0.."""
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = a.copy().reshape(-1, a.shape[-2], a.shape[-1])
        for b in range(out.shape[0]):
            for i in range(out.shape[1]):
                lib().orc_ntt(self.hp, midx0 + i, P(out[b, i]), 1 if inverse else 0)
        return out.reshape(a.shape)

    def basis_extend(self, sidx, src, tidx):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        dst = np.zeros((len(tidx), self.N), dtype=np.uint64)
        s = (C.c_int * len(sidx))(*sidx)
        t = (C.c_int * len(tidx))(*tidx)
        lib().orc_basis_extend(self.hp, s, len(sidx), P(src), t, len(tidx), P(dst))
        return dst

    def permute_ntt_index(self, gal):
        out = np.zeros(self.N, dtype=np.uint32)
        lib().orc_permute_ntt_index(self.p.logN, C.c_uint64(gal), out.ctypes.data_as(C.POINTER(C.c_uint32)))
        return out

    # ---- client-side crypto (test only) ----
    def gen_secret(self, seed):
        s = np.zeros(self.N, dtype=np.int8)
        lib().orc_gen_secret(self.hp, C.c_uint64(seed), s.ctypes.data_as(i8p))
        return s

    def encrypt(self, s, seed, level, pt):
        pt = np.ascontiguousarray(pt, dtype=np.uint64)
        assert pt.shape == (level + 1, self.N)
        ct = np.zeros((2, level + 1, self.N), dtype=np.uint64)
        lib().orc_encrypt_sk(self.hp, s.ctypes.data_as(i8p), C.c_uint64(seed), level, P(pt), P(ct))
        return ct

    def decrypt(self, s, ct):
        ct = np.ascontiguousarray(ct, dtype=np.uint64)
        deg, l = ct.shape[0] - 1, ct.shape[1]
        pt = np.zeros((l, self.N), dtype=np.uint64)
        lib().orc_decrypt(self.hp, s.ctypes.data_as(i8p), l - 1, deg, P(ct), P(pt))
        return pt

    def evk_shape(self, key_level):
        return (self.p.beta(key_level), 2, key_level + 1 + self.p.k, self.N)

    def gen_rlk(self, s, seed, key_level):
        evk = np.zeros(self.evk_shape(key_level), dtype=np.uint64)
        lib().orc_gen_rlk(self.hp, s.ctypes.data_as(i8p), C.c_uint64(seed), key_level, P(evk))
        return evk

    def gen_glk(self, s, seed, key_level, gal):
        evk = np.zeros(self.evk_shape(key_level), dtype=np.uint64)
        lib().orc_gen_glk(self.hp, s.ctypes.data_as(i8p), C.c_uint64(seed), key_level, C.c_uint64(gal), P(evk))
        return evk

    def bfv_scale_up(self, level, m):
        m = np.ascontiguousarray(m, dtype=np.uint64)
        pt = np.zeros((level + 1, self.N), dtype=np.uint64)
        lib().orc_bfv_scale_up(self.hp, level, P(m), P(pt))
        return pt

    # ---- hot-path ops ----
    def keyswitch_ntt(self, c_ntt, evk):
        c_ntt = np.ascontiguousarray(c_ntt, dtype=np.uint64)
        l = c_ntt.shape[0]
        key_level = evk.shape[2] - self.p.k - 1
        p0 = np.zeros((l, self.N), dtype=np.uint64)
        p1 = np.zeros((l, self.N), dtype=np.uint64)
        lib().orc_keyswitch_ntt(self.hp, l - 1, P(c_ntt), P(evk), key_level, P(p0), P(p1))
        return p0, p1

    def add(self, a, b):
        o = np.zeros_like(a)
        lib().orc_add(self.hp, a.shape[1] - 1, a.shape[0], P(a), P(b), P(o))
        return o

    def sub(self, a, b):
        o = np.zeros_like(a)
        lib().orc_sub(self.hp, a.shape[1] - 1, a.shape[0], P(a), P(b), P(o))
        return o

    def neg(self, a):
        o = np.zeros_like(a)
        lib().orc_neg(self.hp, a.shape[1] - 1, a.shape[0], P(a), P(o))
        return o

    def mult(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = np.ascontiguousarray(b, dtype=np.uint64)
        l = a.shape[1]
        o = np.zeros((3, l, self.N), dtype=np.uint64)
        lib().orc_mult(self.hp, l - 1, P(a), P(b), P(o))
        return o

    def relin(self, ct3, rlk):
        ct3 = np.ascontiguousarray(ct3, dtype=np.uint64)
        l = ct3.shape[1]
        key_level = rlk.shape[2] - self.p.k - 1
        o = np.zeros((2, l, self.N), dtype=np.uint64)
        lib().orc_relin(self.hp, l - 1, P(ct3), P(rlk), key_level, P(o))
        return o

    def rescale(self, ct):
        ct = np.ascontiguousarray(ct, dtype=np.uint64)
        npoly, l = ct.shape[0], ct.shape[1]
        o = np.zeros((npoly, l - 1, self.N), dtype=np.uint64)
        lib().orc_rescale(self.hp, l - 1, npoly, P(ct), P(o))
        return o

    def drop_level(self, ct):
        return np.ascontiguousarray(ct[:, :-1, :])

    def rotate(self, ct, glk, gal):
        ct = np.ascontiguousarray(ct, dtype=np.uint64)
        l = ct.shape[1]
        key_level = glk.shape[2] - self.p.k - 1
        o = np.zeros((2, l, self.N), dtype=np.uint64)
        lib().orc_rotate(self.hp, l - 1, P(ct), P(glk), key_level, C.c_uint64(gal), P(o))
        return o


# ---------------------------------------------------------------------------------------------
# Galois elements (frontend/custom_task.py:139-145)
# ---------------------------------------------------------------------------------------------
def galois_for_col_step(N, step):
    return pow(5, step % (2 * N), 2 * N)


def galois_for_row(N):
    return 2 * N - 1


# ---------------------------------------------------------------------------------------------
# CRT helpers (Python integers)
# ---------------------------------------------------------------------------------------------
def crt_centered(limbs, moduli):
    """limbs [l][N] uint64 -> list of centred Python ints mod prod(moduli)."""
    l = len(moduli)
    Q = 1
    for m in moduli:
        Q *= m
    coef = []
    for i, m in enumerate(moduli):
        Qi = Q // m
        coef.append(Qi * pow(Qi % m, -1, m))
    N = limbs.shape[1]
    cols = [limbs[i].tolist() for i in range(l)]
    out = []
    half = Q // 2
    for x in range(N):
        v = 0
        for i in range(l):
            v += cols[i][x] * coef[i]
        v %= Q
        if v > half:
            v -= Q
        out.append(v)
    return out, Q


def ints_to_limbs(vals, moduli):
    """list of Python ints (any sign) -> [l][N] uint64 residues"""
    out = np.zeros((len(moduli), len(vals)), dtype=np.uint64)
    for i, m in enumerate(moduli):
        out[i] = np.array([v % m for v in vals], dtype=np.uint64)
    return out


# ---------------------------------------------------------------------------------------------
# CKKS canonical-embedding encode / decode (slot j <-> evaluation at xi^(5^j), xi = exp(i*pi/N))
# ---------------------------------------------------------------------------------------------
def _rot_group(N):
    g, out = 1, []
    for _ in range(N // 2):
        out.append(g)
        g = g * 5 % (2 * N)
    return np.array(out)


def ckks_encode_coeffs(values, N, scale):
    """complex values (len N/2) -> list of N Python-int coefficients round(scale * m_k)."""
    z = np.asarray(values, dtype=np.complex128)
    assert z.shape[0] == N // 2
    rg = _rot_group(N)
    M = np.zeros(2 * N, dtype=np.complex128)
    M[rg] = z
    M[(2 * N - rg) % (2 * N)] = np.conj(z)
    m = np.fft.fft(M)[:N].real / N
    return [int(round(float(v) * scale)) for v in m]


def ckks_decode_coeffs(coeffs, N, scale):
    m = np.zeros(2 * N, dtype=np.complex128)
    m[:N] = np.array([float(c) for c in coeffs]) / scale
    M = np.fft.ifft(m) * (2 * N)
    return M[_rot_group(N)]


def ckks_encode(orc, values, level, scale):
    """-> NTT-domain plaintext [level+1][N]"""
    coeffs = ckks_encode_coeffs(values, orc.N, scale)
    return orc.ntt_limbs(ints_to_limbs(coeffs, orc.p.q[:level + 1]))


def ckks_decode(orc, pt_ntt, scale):
    l = pt_ntt.shape[0]
    pt = orc.ntt_limbs(pt_ntt, inverse=True)
    use = min(l, 3)
    coeffs, _ = crt_centered(pt[:use], orc.p.q[:use])
    return ckks_decode_coeffs(coeffs, orc.N, scale)


def mean_precision_bits(got, want):
    """mean real precision in bits, as fhe_ops_lib/precision.cpp:103-183 (log2(1/mean |dRe|))"""
    err = np.abs(np.real(np.asarray(got)) - np.real(np.asarray(want)))
    return float(np.log2(1.0 / max(float(np.mean(err)), 2.0 ** -60)))


# ---------------------------------------------------------------------------------------------
# BFV batch encode / decode (Lattigo bfv encoder index matrix; slot i <-> evaluation at psi_t^(5^i))
# ---------------------------------------------------------------------------------------------
class BfvEncoder:
    def __init__(self, params):
        self.N, self.logN, self.t = params.N, params.logN, params.t
        self.ring_t = Oracle(Params("ring_t", SCHEME_CKKS, params.logN, [params.t], []))
        N, logN = self.N, self.logN
        idx = np.zeros(N, dtype=np.int64)
        pos, m = 1, 2 * N
        for i in range(N // 2):
            i1, i2 = (pos - 1) >> 1, (m - pos - 1) >> 1
            idx[i] = int(format(i1, "0%db" % logN)[::-1], 2)
            idx[i | (N >> 1)] = int(format(i2, "0%db" % logN)[::-1], 2)
            pos = pos * 5 % m
        self.index = idx

    def encode_coeffs(self, values):
        """values: N ints mod t (2 rows x N/2) -> coefficient vector mod t"""
        buf = np.zeros(self.N, dtype=np.uint64)
        buf[self.index] = np.asarray(values, dtype=np.uint64) % self.t
        return self.ring_t.ntt(0, buf, inverse=True)

    def decode_coeffs(self, coeffs):
        out = self.ring_t.ntt(0, np.asarray(coeffs, dtype=np.uint64) % self.t)
        return out[self.index]


def bfv_decode_pt(orc, pt):
    """coefficient-domain BFV plaintext [l][N] (Delta*m + noise) -> coefficients mod t (round(t*x/Q) mod t)"""
    l = pt.shape[0]
    t = orc.p.t
    vals, Q = crt_centered(pt, orc.p.q[:l])
    out = np.zeros(orc.N, dtype=np.uint64)
    for i, v in enumerate(vals):
        out[i] = ((2 * t * v + Q) // (2 * Q)) % t
    return out
