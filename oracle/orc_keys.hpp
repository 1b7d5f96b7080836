// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.hpp header).
//
// Test-only client-side crypto so that the hot-path ops can be validated by the reference's own
// criterion (decrypt+decode equals plaintext arithmetic; unittests/test_gpu_bfv.cpp:333-335,
// unittests/test_gpu_ckks.cpp:37-43).  Deterministic from a seed (the reference's helpers are
// random_device-seeded, fhe_ops_lib/utils.cpp:104-107).  Key shapes follow Lattigo v4
// rlwe/keygen.go genSwitchingKey (evk_i = Enc_{s_out}(P * s_in restricted to digit i)) and
// GenSwitchingKeyForGalois (s_out = sigma_{g^-1}(s)), exported in the reference ABI form
// (NTT domain, Montgomery removed; plug-in/lattigo/acc/c_struct_import_export.go:83-135).
#pragma once
#include "orc_core.hpp"

namespace orc {

struct Rng {
    u64 s[4];
    explicit Rng(u64 seed) {
        for (int i = 0; i < 4; i++) {
            seed += 0x9E3779B97F4A7C15ull;
            u64 z = seed;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            s[i] = z ^ (z >> 31);
        }
    }
    static u64 rotl(u64 x, int k) { return (x << k) | (x >> (64 - k)); }
    u64 next() {
        u64 r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    u64 uniform(u64 q) {
        u64 lim = ~0ull - (~0ull % q);
        u64 v;
        do { v = next(); } while (v >= lim);
        return v % q;
    }
    double unit() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
    int gaussian(double sigma = 3.2, int bound = 19) {
        for (;;) {
            double u1 = unit(), u2 = unit();
            if (u1 < 1e-300) continue;
            double g = std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2) * sigma;
            int r = (int)std::llround(g);
            if (r >= -bound && r <= bound) return r;
        }
    }
};

inline void gen_secret(const Context& ctx, u64 seed, int8_t* s) {
    Rng rng(seed);
    for (int x = 0; x < ctx.N; x++) s[x] = (int8_t)((int)rng.uniform(3) - 1);
}

// small signed poly -> NTT form under modulus index midx
inline void small_to_ntt(const Context& ctx, int midx, const int* v, u64* out) {
    u64 q = ctx.mods[midx].q;
    for (int x = 0; x < ctx.N; x++) out[x] = v[x] >= 0 ? (u64)v[x] : q - (u64)(-v[x]);
    ntt_forward(ctx.mods[midx], ctx.N, out);
}
inline void secret_ntt(const Context& ctx, int midx, const int8_t* s, u64* out) {
    std::vector<int> v(ctx.N);
    for (int x = 0; x < ctx.N; x++) v[x] = s[x];
    small_to_ntt(ctx, midx, v.data(), out);
}

// Secret-key encryption of pt ([l][N], native domain of the scheme) -> ct [2][l][N] (native domain).
inline void encrypt_sk(const Context& ctx, const int8_t* s, u64 seed, int level, const u64* pt, u64* ct) {
    int N = ctx.N, l = level + 1;
    Rng rng(seed);
    std::vector<int> e(N);
    for (int x = 0; x < N; x++) e[x] = rng.gaussian();
    std::vector<u64> sn(N), en(N), m(N);
    bool ntt_native = ctx.scheme == SCHEME_CKKS;
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        secret_ntt(ctx, i, s, sn.data());
        small_to_ntt(ctx, i, e.data(), en.data());
        std::memcpy(m.data(), pt + (size_t)i * N, sizeof(u64) * N);
        if (!ntt_native) ntt_forward(ctx.Q(i), N, m.data());
        u64* c0 = ct + (size_t)i * N;
        u64* c1 = ct + (size_t)(l + i) * N;
        for (int x = 0; x < N; x++) {
            u64 a = rng.uniform(q);
            c1[x] = a;
            c0[x] = addmod(submod(m[x], mulmod(a, sn[x], q), q), en[x], q);
        }
        if (!ntt_native) {
            ntt_inverse(ctx.Q(i), N, c0);
            ntt_inverse(ctx.Q(i), N, c1);
        }
    }
}

// pt = sum_d c_d * s^d, native domain in/out.
inline void decrypt(const Context& ctx, const int8_t* s, int level, int degree, const u64* ct, u64* pt) {
    int N = ctx.N, l = level + 1;
    bool ntt_native = ctx.scheme == SCHEME_CKKS;
    std::vector<u64> sn(N), c(N), acc(N);
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        secret_ntt(ctx, i, s, sn.data());
        std::fill(acc.begin(), acc.end(), 0);
        for (int d = degree; d >= 0; d--) {  // Horner
            std::memcpy(c.data(), ct + ((size_t)d * l + i) * N, sizeof(u64) * N);
            if (!ntt_native) ntt_forward(ctx.Q(i), N, c.data());
            for (int x = 0; x < N; x++) acc[x] = addmod(mulmod(acc[x], sn[x], q), c[x], q);
        }
        if (!ntt_native) ntt_inverse(ctx.Q(i), N, acc.data());
        std::memcpy(pt + (size_t)i * N, acc.data(), sizeof(u64) * N);
    }
}

// evk [beta][2][key_level+1+k][N] switching s_in -> s_out, both given in coefficient form (signed small).
// s_in may be a general polynomial given per-modulus in NTT form through the callback.
template <typename SinNtt>
inline void gen_switching_key(const Context& ctx, const int8_t* s_out, SinNtt&& sin_ntt, u64 seed, int key_level,
                              u64* evk) {
    int N = ctx.N, lq = key_level + 1, k = ctx.np, alpha = k;
    int beta = (lq + alpha - 1) / alpha, limbs = lq + k;
    Rng rng(seed);
    std::vector<u64> so(N), si(N), en(N);
    std::vector<int> e(N);
    for (int dg = 0; dg < beta; dg++) {
        for (int x = 0; x < N; x++) e[x] = rng.gaussian();
        for (int tl = 0; tl < limbs; tl++) {
            int midx = tl < lq ? ctx.idxQ(tl) : ctx.idxP(tl - lq);
            u64 q = ctx.mods[midx].q;
            secret_ntt(ctx, midx, s_out, so.data());
            small_to_ntt(ctx, midx, e.data(), en.data());
            u64* b = evk + ((size_t)(dg * 2 + 0) * limbs + tl) * N;
            u64* a = evk + ((size_t)(dg * 2 + 1) * limbs + tl) * N;
            bool in_digit = tl < lq && tl >= dg * alpha && tl < (dg + 1) * alpha;
            u64 pmod = 1;
            if (in_digit) {
                for (int j = 0; j < k; j++) pmod = mulmod(pmod, ctx.P(j).q % q, q);
                sin_ntt(midx, si.data());
            }
            for (int x = 0; x < N; x++) {
                u64 av = rng.uniform(q);
                a[x] = av;
                u64 v = addmod(submod(0, mulmod(av, so[x], q), q), en[x], q);
                if (in_digit) v = addmod(v, mulmod(pmod, si[x], q), q);
                b[x] = v;
            }
        }
    }
}

inline void gen_rlk(const Context& ctx, const int8_t* s, u64 seed, int key_level, u64* evk) {
    gen_switching_key(ctx, s, [&](int midx, u64* out) {
        secret_ntt(ctx, midx, s, out);
        u64 q = ctx.mods[midx].q;
        for (int x = 0; x < ctx.N; x++) out[x] = mulmod(out[x], out[x], q);
    }, seed, key_level, evk);
}

inline void gen_glk(const Context& ctx, const int8_t* s, u64 seed, int key_level, u64 gal, u64* evk) {
    // s_out = sigma_{gal^-1}(s)
    u64 twoN = 2ull << ctx.logN;
    u64 gal_inv = 1;
    {  // gal^(2N-1) mod 2N  (Lattigo: ModExp(galEl, NthRoot-1, NthRoot))
        u64 b = gal % twoN, e = twoN - 1;
        while (e) { if (e & 1) gal_inv = gal_inv * b % twoN; b = b * b % twoN; e >>= 1; }
    }
    std::vector<int8_t> sout(ctx.N);
    u64 mask = ctx.N - 1;
    for (u64 i = 0; i < (u64)ctx.N; i++) {
        u64 index = i * gal_inv;
        bool sign = (index >> ctx.logN) & 1;
        sout[index & mask] = sign ? (int8_t)-s[i] : s[i];
    }
    gen_switching_key(ctx, sout.data(), [&](int midx, u64* out) { secret_ntt(ctx, midx, s, out); }, seed, key_level,
                      evk);
}

// bfv Scaler.ScaleUp (scaleUpTCoprimeWithQVec): pt_i = round(Q_level * m / t) mod q_i, m in [0,t) coefficient form.
inline void bfv_scale_up(const Context& ctx, int level, const u64* m, u64* pt) {
    int N = ctx.N, l = level + 1;
    u64 t = ctx.t, thalf = t >> 1;
    u64 qmodt = 1;
    for (int i = 0; i < l; i++) qmodt = mulmod(qmodt, ctx.Q(i).q % t, t);
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        u64 neg_tinv = negmod(invmod(t % q, q), q);
        u64 thalf_q = thalf % q;
        for (int x = 0; x < N; x++) {
            u64 r = addmod(mulmod(m[x] % t, qmodt, t), thalf, t);  // (m*Q + t/2) mod t
            u64 v = submod(r % q, thalf_q, q);                      // centred
            pt[(size_t)i * N + x] = mulmod(v, neg_tinv, q);
        }
    }
}

}  // namespace orc
