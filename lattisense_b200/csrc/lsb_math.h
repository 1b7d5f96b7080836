// Modular arithmetic shared by the sm_100a kernels (device) and the host-side table builders.
// Everything here is LSB_HD so the same code is exercised on the CPU by tests/emul (index / range
// logic checks without a GPU); the product path only ever runs it inside CUDA kernels.
//
// Conventions: residues are u64 in [0,q), q < 2^62 (the reference's widest prime is 61 bits,
// frontend/custom_task.py:387-420), so lazy ranges up to 4q fit a word.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define LSB_HD __host__ __device__ __forceinline__
#define LSB_D __device__ __forceinline__
#else
#define LSB_HD inline
#define LSB_D inline
#endif

namespace lsb {

typedef uint64_t u64;
typedef uint32_t u32;
typedef unsigned __int128 u128;

// (value, Shoup companion floor(value * 2^64 / q)); 16 bytes so one LDG.128 fetches both.
struct __attribute__((aligned(16))) Tw {
    u64 w;
    u64 ws;
};

LSB_HD u64 umulhi(u64 a, u64 b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (u64)(((u128)a * b) >> 64);
#endif
}

// y * w mod q, lazily: result in [0, 2q) for ANY 64-bit y (Harvey / Shoup).
LSB_HD u64 mul_shoup_lazy(u64 y, Tw t, u64 q) {
    u64 h = umulhi(y, t.ws);
    return y * t.w - h * q;
}
LSB_HD u64 csub(u64 x, u64 m) { return x >= m ? x - m : x; }
LSB_HD u64 mul_shoup(u64 y, Tw t, u64 q) { return csub(mul_shoup_lazy(y, t, q), q); }

LSB_HD u64 addmod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
LSB_HD u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
LSB_HD u64 negmod(u64 a, u64 q) { return a ? q - a : 0; }

// Per-modulus constants for Barrett reduction of 64- and 128-bit values: r = floor(2^128 / q).
struct BarrettQ {
    u64 q;
    u64 r_lo, r_hi;
    // shift-Barrett (reduce_wide): sh = bitlen(q) - 1, mu = floor(2^(sh+64) / q) < 2^64; qinv64 = floor((2^64-1)/q)
    u64 mu, qinv64;
    u32 sh, pad_;
};

// z = (hi:lo) < 2^128, returns z mod q (SEAL-style two-word Barrett; one correction).
LSB_HD u64 barrett128(u64 lo, u64 hi, const BarrettQ& m) {
    // floor(z * r / 2^128) using the three relevant partial products
    u64 c0 = umulhi(lo, m.r_lo);
    u64 t1_lo = lo * m.r_hi, t1_hi = umulhi(lo, m.r_hi);
    u64 s0 = t1_lo + c0;
    u64 carry0 = s0 < c0;
    u64 t2_lo = hi * m.r_lo, t2_hi = umulhi(hi, m.r_lo);
    u64 s1 = s0 + t2_lo;
    u64 carry1 = s1 < t2_lo;
    u64 quot = hi * m.r_hi + t1_hi + t2_hi + carry0 + carry1;
    u64 rem = lo - quot * m.q;
    return csub(rem, m.q);
}
// x = (hi:lo) < 2^(sh+64) -> x mod q with one mul.hi and one mul.lo (the integer multiply pipe is what bounds the
// element-wise kernels on B200: profiles/r01_elem_ncu_summary.txt).  With t = floor(x / 2^sh) < 2^64 and
// qh = floor(t * mu / 2^64):  0 <= x/q - qh < 3, so x - qh*q < 3q < 2^64 and two conditional subtractions finish.
// The bound covers any product of two values below 2q, and sums of n products of residues while n*q < 2^63.
LSB_HD u64 reduce_wide(u64 lo, u64 hi, const BarrettQ& m) {
    const u64 t = (hi << (64 - m.sh)) | (lo >> m.sh);
    const u64 qh = umulhi(t, m.mu);
    u64 r = lo - qh * m.q;
    r = csub(r, m.q << 1);
    return csub(r, m.q);
}
// any 64-bit x -> x mod q
LSB_HD u64 barrett64(u64 x, const BarrettQ& m) {
    u64 r = x - umulhi(x, m.qinv64) * m.q;  // [0, 2q)
    return csub(r, m.q);
}
LSB_HD u64 mulmod(u64 a, u64 b, const BarrettQ& m) { return reduce_wide(a * b, umulhi(a, b), m); }

// 128-bit accumulator for lazy inner products.
struct Acc128 {
    u64 lo, hi;
};
LSB_HD void acc_mad(Acc128& a, u64 x, u64 y) {
    u64 pl = x * y, ph = umulhi(x, y);
    a.lo += pl;
    a.hi += ph + (a.lo < pl);
}

// Forward (Cooley-Tukey, Harvey lazy) butterfly: x in [0,4q), y any -> outputs in [0,4q).
LSB_HD void ct_bfly(u64& x, u64& y, Tw t, u64 q, u64 q2) {
    u64 X = csub(x, q2);
    u64 T = mul_shoup_lazy(y, t, q);
    x = X + T;
    y = X - T + q2;
}
// Same without the range correction on x: every stage adds at most 2q to the bound, so a whole
// logN-stage transform stays below (2 logN + 1) q < 2^64 when q < 2^58 (kLazyLimit).
LSB_HD void ct_bfly_nocorr(u64& x, u64& y, Tw t, u64 q, u64 q2) {
    u64 T = mul_shoup_lazy(y, t, q);
    y = x - T + q2;
    x = x + T;
}
constexpr u64 kLazyLimit = 1ull << 58;
// v < 2^64 -> v mod q, qinv64 = floor((2^64-1)/q)
LSB_HD u64 reduce_full(u64 v, u64 q, u64 qinv64) {
    u64 r = v - umulhi(v, qinv64) * q;  // [0, 2q)
    return csub(r, q);
}

// Inverse (Gentleman-Sande, lazy) butterfly: x,y in [0,2q) -> outputs in [0,2q).
LSB_HD void gs_bfly(u64& x, u64& y, Tw t, u64 q, u64 q2) {
    u64 s = csub(x + y, q2);
    u64 d = x - y + q2;
    y = mul_shoup_lazy(d, t, q);
    x = s;
}

}  // namespace lsb
