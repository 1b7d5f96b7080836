// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.hpp header).  PARITY UNPINNED.
//
// OPTIMISED CPU baseline for bench.py's reference arm: the same algorithms as orc_core.hpp (which restates the
// reference's CPU path, mega_ag_runners/cpu/mega_ag_executors_cpu.cpp:186-295 -> Lattigo v4), written the way a
// production CPU library writes them — Harvey/Shoup lazy butterflies (Lattigo ring/ntt.go keeps values in [0, 2q) /
// [0, 4q) the same way), precomputed Shoup constants for every fixed multiplier, 128-bit lazy accumulation with one
// Barrett reduction per coefficient in the CRT sums and in the key-switch inner product (Lattigo's
// MulCoeffsMontgomeryConstantAndAddNoMod idea), no division in any inner loop, scratch allocated once.
// tests/test_oracle_cpu.py asserts that every output limb equals the simple oracle's.  Only CKKS mult / relin /
// rescale / rotate are provided: the operations bench.py times.
#pragma once
#include <cstring>
#include <memory>
#include <vector>
#include "orc_core.hpp"

namespace orcf {

using orc::u64;
using orc::u128;

struct FMod {
    u64 q = 0, q2 = 0;
    u64 mu_hi = 0, mu_lo = 0;  // floor(2^128 / q)
    std::vector<u64> psi, psi_s, ipsi, ipsi_s;  // bit-reversed powers + Shoup companions
    u64 ninv = 0, ninv_s = 0;
};

static inline u64 shoup_of(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
// w * y mod q in [0, 2q) for any y < 2^64
static inline u64 mul_shoup_lazy(u64 y, u64 w, u64 ws, u64 q) {
    const u64 hi = (u64)(((u128)y * ws) >> 64);
    return w * y - hi * q;
}
static inline u64 mul_shoup(u64 y, u64 w, u64 ws, u64 q) {
    const u64 r = mul_shoup_lazy(y, w, ws, q);
    return r >= q ? r - q : r;
}
// x mod q for x < 2^126 (q < 2^62): quotient estimate from the high words of x * mu, at most 3 corrections
static inline u64 barrett128(u128 x, const FMod& m) {
    const u64 xh = (u64)(x >> 64), xl = (u64)x;
    const u64 qhat = xh * m.mu_hi + (u64)(((u128)xh * m.mu_lo) >> 64) + (u64)(((u128)xl * m.mu_hi) >> 64);
    u64 r = xl - qhat * m.q;
    while (r >= m.q) r -= m.q;
    return r;
}
static inline u64 mulmod_b(u64 a, u64 b, const FMod& m) { return barrett128((u128)a * b, m); }

inline FMod make_fmod(const orc::Modulus& o) {
    FMod m;
    m.q = o.q;
    m.q2 = o.q << 1;
    const u128 mu = ~(u128)0 / o.q;  // floor((2^128 - 1) / q) == floor(2^128 / q) for odd q > 1
    m.mu_hi = (u64)(mu >> 64);
    m.mu_lo = (u64)mu;
    const size_t N = o.psi_br.size();
    m.psi = o.psi_br;
    m.ipsi = o.ipsi_br;
    m.psi_s.resize(N);
    m.ipsi_s.resize(N);
    for (size_t i = 0; i < N; i++) {
        m.psi_s[i] = shoup_of(m.psi[i], o.q);
        m.ipsi_s[i] = shoup_of(m.ipsi[i], o.q);
    }
    m.ninv = o.n_inv;
    m.ninv_s = shoup_of(o.n_inv, o.q);
    return m;
}

// forward: inputs in [0, q), values kept in [0, 4q), outputs reduced to [0, q)
inline void ntt_forward(const FMod& m, int N, u64* a) {
    const u64 q = m.q, q2 = m.q2;
    int t = N;
    for (int mm = 1; mm < N; mm <<= 1) {
        t >>= 1;
        for (int i = 0; i < mm; i++) {
            const u64 w = m.psi[mm + i], ws = m.psi_s[mm + i];
            u64* x = a + 2 * i * t;
            u64* y = x + t;
            for (int j = 0; j < t; j++) {
                u64 U = x[j];
                U = U >= q2 ? U - q2 : U;
                const u64 V = mul_shoup_lazy(y[j], w, ws, q);
                x[j] = U + V;
                y[j] = U - V + q2;
            }
        }
    }
    for (int j = 0; j < N; j++) {
        u64 v = a[j];
        v = v >= q2 ? v - q2 : v;
        a[j] = v >= q ? v - q : v;
    }
}
// inverse: values kept in [0, 2q), N^-1 folded at the end, outputs in [0, q)
inline void ntt_inverse(const FMod& m, int N, u64* a) {
    const u64 q = m.q, q2 = m.q2;
    int t = 1;
    for (int mm = N; mm > 1; mm >>= 1) {
        const int h = mm >> 1;
        for (int i = 0; i < h; i++) {
            const u64 w = m.ipsi[h + i], ws = m.ipsi_s[h + i];
            u64* x = a + 2 * i * t;
            u64* y = x + t;
            for (int j = 0; j < t; j++) {
                const u64 U = x[j], V = y[j];
                u64 S = U + V;
                S = S >= q2 ? S - q2 : S;
                x[j] = S;
                y[j] = mul_shoup_lazy(U - V + q2, w, ws, q);
            }
        }
        t <<= 1;
    }
    for (int j = 0; j < N; j++) a[j] = mul_shoup(a[j], m.ninv, m.ninv_s, q);
}

// CRT tables of one basis extension (same quantities as orc::basis_extend computes per call)
struct BexTab {
    std::vector<int> sidx, tidx;
    std::vector<u64> qhat_inv, qhat_inv_s;        // [ns]
    std::vector<double> s_f64;                    // [ns]
    std::vector<std::vector<u64>> qhat_mod;       // [nt][ns]
    std::vector<std::vector<u64>> neg_vS;         // [nt][ns+1]: p_t - (v * S mod p_t)
};

struct FContext {
    const orc::Context* base = nullptr;
    int N = 0, nq = 0, np = 0;
    std::vector<FMod> mods;                              // Q..., P...
    std::vector<std::vector<BexTab>> modup;              // [level][digit]
    std::vector<BexTab> moddown;                         // [level]: P -> Q_0..Q_level
    std::vector<std::vector<u64>> pinv, pinv_s;          // [level][i]: P^-1 mod q_i
    std::vector<std::vector<u64>> qlinv, qlinv_s;        // [level][i]: q_level^-1 mod q_i
    // scratch (one context per thread)
    std::vector<u64> c_coeff, d, ext, last, tmp, p0, p1, ct3, ct2;
    std::vector<u128> acc0, acc1;
    std::vector<u64> y;
};

inline BexTab make_bex(const FContext& c, const std::vector<int>& sidx, const std::vector<int>& tidx) {
    BexTab t;
    t.sidx = sidx;
    t.tidx = tidx;
    const int ns = (int)sidx.size(), nt = (int)tidx.size();
    t.qhat_inv.resize(ns);
    t.qhat_inv_s.resize(ns);
    t.s_f64.resize(ns);
    for (int i = 0; i < ns; i++) {
        const u64 si = c.mods[sidx[i]].q;
        u64 prod = 1;
        for (int k = 0; k < ns; k++)
            if (k != i) prod = orc::mulmod(prod, c.mods[sidx[k]].q % si, si);
        t.qhat_inv[i] = orc::invmod(prod, si);
        t.qhat_inv_s[i] = shoup_of(t.qhat_inv[i], si);
        t.s_f64[i] = (double)si;
    }
    t.qhat_mod.assign(nt, std::vector<u64>(ns));
    t.neg_vS.assign(nt, std::vector<u64>(ns + 1));
    for (int j = 0; j < nt; j++) {
        const u64 pj = c.mods[tidx[j]].q;
        u64 all = 1;
        for (int k = 0; k < ns; k++) all = orc::mulmod(all, c.mods[sidx[k]].q % pj, pj);
        for (int v = 0; v <= ns; v++) t.neg_vS[j][v] = orc::negmod(orc::mulmod((u64)v % pj, all, pj), pj);
        for (int i = 0; i < ns; i++) {
            u64 prod = 1;
            for (int k = 0; k < ns; k++)
                if (k != i) prod = orc::mulmod(prod, c.mods[sidx[k]].q % pj, pj);
            t.qhat_mod[j][i] = prod;
        }
    }
    return t;
}

inline std::unique_ptr<FContext> make_fcontext(const orc::Context& o) {
    auto c = std::make_unique<FContext>();
    c->base = &o;
    c->N = o.N;
    c->nq = o.nq;
    c->np = o.np;
    for (int i = 0; i < o.nq + o.np; i++) c->mods.push_back(make_fmod(o.mods[i]));
    const int k = o.np;
    c->modup.resize(o.nq);
    c->moddown.resize(o.nq);
    c->pinv.resize(o.nq);
    c->pinv_s.resize(o.nq);
    c->qlinv.resize(o.nq);
    c->qlinv_s.resize(o.nq);
    for (int level = 0; level < o.nq; level++) {
        const int l = level + 1, beta = (l + k - 1) / k;
        for (int dg = 0; dg < beta; dg++) {
            const int st = dg * k, ed = std::min(st + k, l);
            std::vector<int> s, t;
            for (int i = st; i < ed; i++) s.push_back(i);
            for (int i = 0; i < l; i++)
                if (i < st || i >= ed) t.push_back(i);
            for (int j = 0; j < k; j++) t.push_back(o.nq + j);
            c->modup[level].push_back(make_bex(*c, s, t));
        }
        std::vector<int> s, t;
        for (int j = 0; j < k; j++) s.push_back(o.nq + j);
        for (int i = 0; i < l; i++) t.push_back(i);
        c->moddown[level] = make_bex(*c, s, t);
        for (int i = 0; i < l; i++) {
            const u64 q = o.mods[i].q;
            u64 pv = 1;
            for (int j = 0; j < k; j++) pv = orc::mulmod(pv, o.mods[o.nq + j].q % q, q);
            pv = orc::invmod(pv, q);
            c->pinv[level].push_back(pv);
            c->pinv_s[level].push_back(shoup_of(pv, q));
            const u64 ql = o.mods[level].q;
            const u64 qi = i < level ? orc::invmod(ql % q, q) : 0;
            c->qlinv[level].push_back(qi);
            c->qlinv_s[level].push_back(i < level ? shoup_of(qi, q) : 0);
        }
    }
    const size_t N = o.N, lk = o.nq + o.np;
    c->c_coeff.resize(o.nq * N);
    c->d.resize(lk * N);
    c->ext.resize(o.nq * N);
    c->last.resize(N);
    c->tmp.resize(N);
    c->p0.resize(o.nq * N);
    c->p1.resize(o.nq * N);
    c->ct3.resize(3 * o.nq * N);
    c->ct2.resize(2 * o.nq * N);
    c->acc0.resize(lk * N);
    c->acc1.resize(lk * N);
    c->y.resize(16);
    return c;
}

// orc::basis_extend with the tables hoisted; src[i] / dst[j] are limb pointers
inline void basis_extend(const FContext& c, const BexTab& t, const u64* const* src, u64* const* dst) {
    const int ns = (int)t.sidx.size(), nt = (int)t.tidx.size(), N = c.N;
    u64 y[16];
    for (int x = 0; x < N; x++) {
        double vf = 0.0;
        for (int i = 0; i < ns; i++) {
            const FMod& m = c.mods[t.sidx[i]];
            y[i] = mul_shoup(src[i][x], t.qhat_inv[i], t.qhat_inv_s[i], m.q);
            vf += (double)y[i] / t.s_f64[i];
        }
        const u64 v = (u64)vf;
        for (int j = 0; j < nt; j++) {
            const FMod& m = c.mods[t.tidx[j]];
            u128 acc = t.neg_vS[j][v];
            for (int i = 0; i < ns; i++) acc += (u128)y[i] * t.qhat_mod[j][i];
            dst[j][x] = barrett128(acc, m);
        }
    }
}
inline void centered_lift(const FContext& c, int sidx, const u64* src, const std::vector<int>& tidx, u64* const* dst) {
    const int N = c.N, nt = (int)tidx.size();
    const u64 s = c.mods[sidx].q;
    for (int x = 0; x < N; x++) {
        u64 v = src[x];
        const bool neg = v >= (s >> 1);
        if (neg) v = s - v;
        for (int j = 0; j < nt; j++) {
            const FMod& m = c.mods[tidx[j]];
            const u64 r = barrett128(v, m);
            dst[j][x] = neg ? orc::negmod(r, m.q) : r;
        }
    }
}

// orc::keyswitch_ntt, optimised
inline void keyswitch_ntt(FContext& c, int level, const u64* c_ntt, const u64* evk, int key_level, u64* p0, u64* p1) {
    const int N = c.N, l = level + 1, k = c.np, alpha = k, lk = l + k;
    const int beta = (l + alpha - 1) / alpha, klimbs = key_level + 1 + k;
    std::memcpy(c.c_coeff.data(), c_ntt, sizeof(u64) * l * N);
    for (int i = 0; i < l; i++) ntt_inverse(c.mods[i], N, &c.c_coeff[(size_t)i * N]);
    std::fill(c.acc0.begin(), c.acc0.begin() + (size_t)lk * N, (u128)0);
    std::fill(c.acc1.begin(), c.acc1.begin() + (size_t)lk * N, (u128)0);
    for (int dg = 0; dg < beta; dg++) {
        const BexTab& t = c.modup[level][dg];
        const int st = dg * alpha, ed = std::min(st + alpha, l);
        const u64* src[16];
        u64* dst[64];
        for (int i = st; i < ed; i++) src[i - st] = &c.c_coeff[(size_t)i * N];
        for (size_t j = 0; j < t.tidx.size(); j++) {
            const int ti = t.tidx[j];
            dst[j] = &c.d[(size_t)(ti < c.nq ? ti : l + (ti - c.nq)) * N];
        }
        if (ed - st == 1) centered_lift(c, st, src[0], t.tidx, dst);
        else basis_extend(c, t, src, dst);
        for (size_t j = 0; j < t.tidx.size(); j++) ntt_forward(c.mods[t.tidx[j]], N, dst[j]);
        for (int tl = 0; tl < lk; tl++) {
            const int key_limb = tl < l ? tl : key_level + 1 + (tl - l);
            const bool own = tl >= st && tl < ed;
            const u64* dd = own ? c_ntt + (size_t)tl * N : &c.d[(size_t)tl * N];
            const u64* e0 = evk + ((size_t)(dg * 2 + 0) * klimbs + key_limb) * N;
            const u64* e1 = evk + ((size_t)(dg * 2 + 1) * klimbs + key_limb) * N;
            u128* a0 = &c.acc0[(size_t)tl * N];
            u128* a1 = &c.acc1[(size_t)tl * N];
            // q < 2^62: a product is below 2^124, so up to 16 digits accumulate without reduction (beta <= 5 here)
            for (int x = 0; x < N; x++) {
                a0[x] += (u128)dd[x] * e0[x];
                a1[x] += (u128)dd[x] * e1[x];
            }
        }
    }
    for (int b = 0; b < 2; b++) {
        std::vector<u128>& acc = b ? c.acc1 : c.acc0;
        u64* out = b ? p1 : p0;
        const BexTab& t = c.moddown[level];
        const u64* src[16];
        u64* dst[64];
        // P limbs of the accumulator -> coefficient domain (reuse c.d as scratch)
        for (int j = 0; j < k; j++) {
            u64* pj = &c.d[(size_t)j * N];
            const FMod& m = c.mods[c.nq + j];
            const u128* a = &acc[(size_t)(l + j) * N];
            for (int x = 0; x < N; x++) pj[x] = barrett128(a[x], m);
            ntt_inverse(m, N, pj);
            src[j] = pj;
        }
        for (int i = 0; i < l; i++) dst[i] = &c.ext[(size_t)i * N];
        basis_extend(c, t, src, dst);
        for (int i = 0; i < l; i++) {
            const FMod& m = c.mods[i];
            ntt_forward(m, N, dst[i]);
            const u64 pv = c.pinv[level][i], pvs = c.pinv_s[level][i];
            const u128* a = &acc[(size_t)i * N];
            for (int x = 0; x < N; x++)
                out[(size_t)i * N + x] = mul_shoup(orc::submod(barrett128(a[x], m), dst[i][x], m.q), pv, pvs, m.q);
        }
    }
}

inline void ckks_mult(FContext& c, int level, const u64* a, const u64* b, u64* o) {
    const int N = c.N, l = level + 1;
    for (int i = 0; i < l; i++) {
        const FMod& m = c.mods[i];
        const u64 *a0 = a + (size_t)i * N, *a1 = a + (size_t)(l + i) * N;
        const u64 *b0 = b + (size_t)i * N, *b1 = b + (size_t)(l + i) * N;
        u64 *o0 = o + (size_t)i * N, *o1 = o + (size_t)(l + i) * N, *o2 = o + (size_t)(2 * l + i) * N;
        for (int x = 0; x < N; x++) {
            o0[x] = mulmod_b(a0[x], b0[x], m);
            o1[x] = barrett128((u128)a0[x] * b1[x] + (u128)a1[x] * b0[x], m);
            o2[x] = mulmod_b(a1[x], b1[x], m);
        }
    }
}
inline void ckks_relin(FContext& c, int level, const u64* ct3, const u64* rlk, int key_level, u64* out2) {
    const int N = c.N, l = level + 1;
    keyswitch_ntt(c, level, ct3 + (size_t)2 * l * N, rlk, key_level, c.p0.data(), c.p1.data());
    for (int i = 0; i < l; i++) {
        const u64 q = c.mods[i].q;
        for (int x = 0; x < N; x++) {
            out2[(size_t)i * N + x] = orc::addmod(ct3[(size_t)i * N + x], c.p0[(size_t)i * N + x], q);
            out2[(size_t)(l + i) * N + x] = orc::addmod(ct3[(size_t)(l + i) * N + x], c.p1[(size_t)i * N + x], q);
        }
    }
}
inline void ckks_rescale(FContext& c, int level, int npoly, const u64* in, u64* out) {
    const int N = c.N, l = level + 1;
    const FMod& ml = c.mods[level];
    const u64 ql = ml.q, half = (ql - 1) >> 1;
    for (int p = 0; p < npoly; p++) {
        std::memcpy(c.last.data(), in + ((size_t)p * l + level) * N, sizeof(u64) * N);
        ntt_inverse(ml, N, c.last.data());
        for (int x = 0; x < N; x++) c.last[x] = orc::addmod(c.last[x], half, ql);
        for (int i = 0; i < level; i++) {
            const FMod& m = c.mods[i];
            const u64 half_q = half % m.q;
            for (int x = 0; x < N; x++) c.tmp[x] = orc::submod(barrett128(c.last[x], m), half_q, m.q);
            ntt_forward(m, N, c.tmp.data());
            const u64* src = in + ((size_t)p * l + i) * N;
            u64* dst = out + ((size_t)p * (l - 1) + i) * N;
            const u64 qi = c.qlinv[level][i], qis = c.qlinv_s[level][i];
            for (int x = 0; x < N; x++) dst[x] = mul_shoup(orc::submod(src[x], c.tmp[x], m.q), qi, qis, m.q);
        }
    }
}
inline void ckks_mult_relin_rescale(FContext& c, int level, const u64* a, const u64* b, const u64* rlk, int key_level,
                                    u64* out) {
    ckks_mult(c, level, a, b, c.ct3.data());
    ckks_relin(c, level, c.ct3.data(), rlk, key_level, c.ct2.data());
    ckks_rescale(c, level, 2, c.ct2.data(), out);
}
inline void ckks_rotate(FContext& c, int level, const u64* ct, const u64* glk, int key_level, const uint32_t* idx,
                        u64* out) {
    const int N = c.N, l = level + 1;
    keyswitch_ntt(c, level, ct + (size_t)l * N, glk, key_level, c.p0.data(), c.p1.data());
    for (int i = 0; i < l; i++) {
        const u64 q = c.mods[i].q;
        u64* a = &c.p0[(size_t)i * N];
        for (int x = 0; x < N; x++) a[x] = orc::addmod(a[x], ct[(size_t)i * N + x], q);
        for (int x = 0; x < N; x++) out[(size_t)i * N + x] = a[idx[x]];
        for (int x = 0; x < N; x++) out[(size_t)(l + i) * N + x] = c.p1[(size_t)i * N + idx[x]];
    }
}

}  // namespace orcf
