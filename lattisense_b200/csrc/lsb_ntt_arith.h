// Butterfly arithmetic policies of the cluster NTT (lsb_ntt_cluster.cu).
//
//  * ArithU64 — Harvey/Shoup lazy butterflies on 64-bit residues (any prime < 2^62).  On B200 this path is
//    bound by the quarter-rate IMAD.WIDE (3.4 butterflies/clk/SM, profiles/r01_int_pipe_microbench.txt).
//  * ArithF64 — the same butterflies on exact integers held in IEEE doubles, for primes with
//    q * (logN + 1) < 2^51 (every default CKKS prime).  y*w mod q is computed from the error-free product
//    (h, l) = two_prod(y, w), the quotient estimate rint(y * (w/q)) (one FMA with the 1.5*2^52 magic constant)
//    and r = fma(-qf, q, h) + l, all of which are exact for |y| < 2^51, q < 2^48: 8 FP64-pipe instructions
//    per butterfly, 7.7 butterflies/clk/SM.  |r| <= 0.75 q, so a forward transform grows by < q per stage and
//    never needs a correction; the inverse (sum path doubles per stage) is re-centred once per round when
//    q * 2^logN >= 2^51.  Outputs are reduced to the canonical residue, so both policies are bit-identical.
//
// Everything is LSB_HD: tests/emul runs the same code on the CPU.
#pragma once
#include <cmath>
#include "lsb_math.h"

namespace lsb {

struct __attribute__((aligned(16))) TwD {
    double w;   // twiddle as a double
    double wq;  // w / q, correctly rounded
};

// per-limb constants handed to the butterflies (CTA-uniform)
struct ModCtx {
    u64 q, q2, qinv64;
    double qd, qinvd;
    bool lazy;    // integer path: q < 2^58, uncorrected forward butterflies
    bool recentre;  // FP64 inverse: reduce after every round
};

constexpr double kMagic = 6755399441055744.0;  // 1.5 * 2^52
constexpr double kTwo52 = 4503599627370496.0;

LSB_HD double d_mul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    return a * b;
#endif
}
LSB_HD double d_add(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
}
LSB_HD double d_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return std::fma(a, b, c);
#endif
}
LSB_HD double bits_to_double(u64 b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    double d;
    __builtin_memcpy(&d, &b, 8);
    return d;
#endif
}
LSB_HD u64 double_to_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (u64)__double_as_longlong(d);
#else
    u64 b;
    __builtin_memcpy(&b, &d, 8);
    return b;
#endif
}

// y * w mod q, result in [-0.75 q, 0.75 q] for |y| < 2^51
LSB_HD double f64_mulmod(double y, TwD t, double q) {
    const double h = d_mul(y, t.w);
    const double l = d_fma(y, t.w, -h);
    // (FRND.F64 in place of the magic-constant pair was measured: no gain, 2561 vs 2596 GB/s at N=2^14)
    const double qf = d_add(d_fma(y, t.wq, kMagic), -kMagic);
    return d_add(d_fma(-qf, q, h), l);
}
// v -> v mod q in [-q/2, q/2] (|v| < 2^51)
LSB_HD double f64_centre(double v, double q, double qinv) {
    const double qf = d_add(d_fma(v, qinv, kMagic), -kMagic);
    return d_fma(-qf, q, v);
}

// ---- FP64 modular arithmetic for the element-wise kernels (primes below 2^47, same exactness argument) ----
// a * b mod q for two variable operands |a|, |b| < 2^48 (no precomputed b/q): the quotient estimate rint(h / q) may be
// off by one, so |result| <= 1.5 q; h - qf*q is an integer below 2^50 and therefore exact in the FMA.
LSB_HD double f64_mulmod_var(double a, double b, double q, double qinv) {
    const double h = d_mul(a, b);
    const double l = d_fma(a, b, -h);
    const double qf = d_add(d_fma(h, qinv, kMagic), -kMagic);
    return d_add(d_fma(-qf, q, h), l);
}
LSB_HD double f64_from_u64(u64 x) { return d_add(bits_to_double(x | 0x4330000000000000ull), -kTwo52); }  // x < 2^52
// |v| < 2^51 -> canonical residue in [0, q) as a double / as an integer
LSB_HD double f64_canon_d(double v, double q, double qinv) {
    const double r = f64_centre(v, q, qinv);
    return r < 0.0 ? d_add(r, q) : r;
}
LSB_HD u64 f64_canon_u(double v, double q, double qinv) {
    const double r = f64_centre(v, q, qinv);
    const double c = r < 0.0 ? d_add(q, kTwo52) : kTwo52;
    return double_to_bits(d_add(r, c)) & 0x000FFFFFFFFFFFFFull;
}
// primes the element-wise FP64 paths accept (independent of the ring degree)
constexpr u64 kElemF64Limit = 1ull << 47;

struct ArithU64 {
    typedef u64 V;
    typedef Tw TW;
    static LSB_HD V from_raw(u64 x) { return x; }
    static LSB_HD V from_bits(u64 b) { return b; }
    static LSB_HD u64 to_bits(V v) { return v; }
    static LSB_HD void ct(V& x, V& y, TW t, const ModCtx& m) {
        if (m.lazy) ct_bfly_nocorr(x, y, t, m.q, m.q2);
        else ct_bfly(x, y, t, m.q, m.q2);
    }
    static LSB_HD void gs(V& x, V& y, TW t, const ModCtx& m) { gs_bfly(x, y, t, m.q, m.q2); }
    // last inverse stage with N^-1 folded: x' = (x+y) * ninv, y' = (x-y) * w_ninv, canonical
    static LSB_HD void gs_last(V& x, V& y, TW ninv, TW w_ninv, const ModCtx& m) {
        const u64 U = x, W = y;
        x = mul_shoup(U + W, ninv, m.q);
        y = mul_shoup(U - W + m.q2, w_ninv, m.q);
    }
    static LSB_HD void end_of_inverse_round(V&, const ModCtx&) {}
    static LSB_HD u64 fwd_canonical(V v, const ModCtx& m) {
        return m.lazy ? reduce_full(v, m.q, m.qinv64) : csub(csub(v, m.q2), m.q);
    }
    static LSB_HD u64 inv_canonical(V v, const ModCtx&) { return v; }  // gs_last already reduced
};

struct ArithF64 {
    typedef double V;
    typedef TwD TW;
    static LSB_HD V from_raw(u64 x) { return d_add(bits_to_double(x | 0x4330000000000000ull), -kTwo52); }  // x < 2^52
    static LSB_HD V from_bits(u64 b) { return bits_to_double(b); }
    static LSB_HD u64 to_bits(V v) { return double_to_bits(v); }
    static LSB_HD void ct(V& x, V& y, TW t, const ModCtx& m) {
        const double r = f64_mulmod(y, t, m.qd);
        y = d_add(x, -r);
        x = d_add(x, r);
    }
    static LSB_HD void gs(V& x, V& y, TW t, const ModCtx& m) {
        const double s = d_add(x, y), d = d_add(x, -y);
        y = f64_mulmod(d, t, m.qd);
        x = s;
    }
    static LSB_HD void gs_last(V& x, V& y, TW ninv, TW w_ninv, const ModCtx& m) {
        const double s = d_add(x, y), d = d_add(x, -y);
        x = f64_mulmod(s, ninv, m.qd);
        y = f64_mulmod(d, w_ninv, m.qd);
    }
    static LSB_HD void end_of_inverse_round(V& v, const ModCtx& m) {
        if (m.recentre) v = f64_centre(v, m.qd, m.qinvd);
    }
    // |v| < 2^51 -> canonical residue in [0, q) as an integer
    static LSB_HD u64 canonical(V v, const ModCtx& m) {
        const double r = f64_centre(v, m.qd, m.qinvd);
        const double c = r < 0.0 ? d_add(m.qd, kTwo52) : kTwo52;
        return double_to_bits(d_add(r, c)) & 0x000FFFFFFFFFFFFFull;
    }
    static LSB_HD u64 fwd_canonical(V v, const ModCtx& m) { return canonical(v, m); }
    static LSB_HD u64 inv_canonical(V v, const ModCtx& m) { return canonical(v, m); }
};

// ArithF64 with 8-byte twiddles: the table holds w only and the quotient constant is formed where the twiddle is loaded,
// wq = fl(w * fl(1/q)) -- one extra DMUL per twiddle (shared by the butterflies that use it) for half the twiddle bytes
// through L1/L2, the unit the cluster kernel keeps busiest (l1tex 77 %, profiles/r02_ntt14_cluster_ncu.txt).  wq is within
// 2^-52 (relative) of w/q instead of 2^-53, so for |y| < 2^51 the quotient estimate is off by at most 1 (0.5 from the
// constant, 0.5 from rint) and |y*w mod q| <= q instead of 0.75 q: still "less than q per stage", which is what the
// eligibility bound q*(logN+1) < 2^51 assumes; h - qf*q stays an integer below 2^50, exact in the FMA.  Outputs are
// reduced to the canonical residue, so the limbs are identical to both other policies (tests/emul).
struct ArithF64W : ArithF64 {
    typedef double TW;
    static LSB_HD double mulmod(double y, double w, const ModCtx& m) {
        const double wq = d_mul(w, m.qinvd);
        const double h = d_mul(y, w);
        const double l = d_fma(y, w, -h);
        const double qf = d_add(d_fma(y, wq, kMagic), -kMagic);
        return d_add(d_fma(-qf, m.qd, h), l);
    }
    static LSB_HD void ct(V& x, V& y, TW t, const ModCtx& m) {
        const double r = mulmod(y, t, m);
        y = d_add(x, -r);
        x = d_add(x, r);
    }
    static LSB_HD void gs(V& x, V& y, TW t, const ModCtx& m) {
        const double s = d_add(x, y), d = d_add(x, -y);
        y = mulmod(d, t, m);
        x = s;
    }
    static LSB_HD void gs_last(V& x, V& y, TW ninv, TW w_ninv, const ModCtx& m) {
        const double s = d_add(x, y), d = d_add(x, -y);
        x = mulmod(s, ninv, m);
        y = mulmod(d, w_ninv, m);
    }
};

// LR consecutive forward stages on R = 2^LR values; tbase = twiddle index of the round's first stage for
// this group (see lsb_ntt_core.h)
//
// S1 >= 0 selects the "tail-SoA" table layout (host::soa_tail in lsb_tables.h) used by the contiguous last round of
// the cluster NTT, whose first global stage is S1: there every thread has its own tbase, and in the standard
// bit-reversed layout a warp's load of twiddle (r, kg) touches addresses 2^r * 16 bytes apart (up to 32
// wavefronts per load).  In the SoA layout entry (r, kg) of group tl = tbase - 2^S1 lives at
// ((2^r | kg) << S1) + tl, so consecutive lanes read consecutive 16-byte entries (4 wavefronts per load).
template <int S1>
LSB_HD u32 tw_index(u32 tbase, int r, int kg) {
    if (S1 >= 0) return ((((u32)1 << r) | (u32)kg) << (S1 >= 0 ? S1 : 0)) + (tbase - ((u32)1 << (S1 >= 0 ? S1 : 0)));
    return (tbase << r) | (u32)kg;
}
template <class P, int LR, int S1 = -1>
LSB_HD void fwd_stages(typename P::V (&x)[1 << LR], const typename P::TW* __restrict__ tw, u32 tbase, const ModCtx& m) {
    constexpr int R = 1 << LR;
#pragma unroll
    for (int r = 0; r < LR; r++) {
        const int hk = R >> (r + 1);
#pragma unroll
        for (int kg = 0; kg < (1 << r); kg++) {
            const typename P::TW t = tw[tw_index<S1>(tbase, r, kg)];
#pragma unroll
            for (int i = 0; i < hk; i++) P::ct(x[kg * 2 * hk + i], x[kg * 2 * hk + i + hk], t, m);
        }
    }
}
// the same stages inverted (Gentleman-Sande, last stage first).  FINAL: the round contains global stage 0.
// RHI / RLO: only the stages r = RHI .. RLO of the round (the whole round by default) -- the cluster kernel's inverse cross
// round runs its first stage, signals the cluster barrier, then runs the rest
template <class P, int LR, bool FINAL, int S1 = -1, int RHI = LR - 1, int RLO = 0>
LSB_HD void inv_stages(typename P::V (&x)[1 << LR], const typename P::TW* __restrict__ tw, u32 tbase, const ModCtx& m,
                       typename P::TW ninv, typename P::TW w_ninv) {
    constexpr int R = 1 << LR;
#pragma unroll
    for (int r = RHI; r >= RLO; r--) {
        const int hk = R >> (r + 1);
        if (FINAL && r == 0) {
#pragma unroll
            for (int i = 0; i < hk; i++) P::gs_last(x[i], x[i + hk], ninv, w_ninv, m);
        } else {
#pragma unroll
            for (int kg = 0; kg < (1 << r); kg++) {
                const typename P::TW t = tw[tw_index<S1>(tbase, r, kg)];
#pragma unroll
                for (int i = 0; i < hk; i++) P::gs(x[kg * 2 * hk + i], x[kg * 2 * hk + i + hk], t, m);
            }
        }
    }
    if (!FINAL && RLO == 0) {
#pragma unroll
        for (int k = 0; k < R; k++) P::end_of_inverse_round(x[k], m);
    }
}

// eligibility of a prime for the FP64 policy
inline bool f64_eligible(u64 q, int logN) { return (unsigned __int128)q * (unsigned)(logN + 1) < ((unsigned __int128)1 << 51); }
inline bool f64_needs_recentre(u64 q, int logN) { return ((unsigned __int128)q << logN) >= ((unsigned __int128)1 << 51); }

}  // namespace lsb
