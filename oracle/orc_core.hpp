// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.hpp header).  PARITY UNPINNED.
//
// CPU restatement of the arithmetic the reference's CPU path (FheTaskCpu) delegates to Lattigo:
//   mega_ag_runners/cpu/mega_ag_executors_cpu.cpp:186,230 (mult)  :236-242 (relin)  :244-255 (rescale)
//   :257-267 (drop_level)  :269-283 (rotate_col)  :285-295 (rotate_row / conjugate)
//   fhe_ops_lib/fhe_lib_v2.cpp:380,408,419,446 (Bfv*)  :970,986,994,1005 (Ckks*)
// The algorithms themselves are the published Lattigo v4 ones (module absent from /root/reference):
//   ring/ntt.go (CT forward / GS inverse, psi powers in bit-reversed order),
//   ring/basis_extension.go (ModUp with float64 overflow count, ModDownQPtoQNTT, ModDownQPtoP),
//   rlwe/keyswitch.go (DecomposeNTT / DecomposeAndSplit / gadget product),
//   ckks/evaluator.go (tensor, DivRoundByLastModulusNTT, permuteNTT),
//   bfv/evaluator.go (tensorAndRescale / quantize), ring/automorphism.go.
// Layout conventions follow the reference ABI (abi/c_types.h:26-60, plug-in/lattigo/acc/
// c_struct_import_export.go:41-59): residues in [0,q), limbs q_0..q_level then (keys) p_0..p_{k-1}.
#pragma once
#include <cmath>
#include <cstring>
#include <memory>
#include "orc_math.hpp"

namespace orc {

enum Scheme { SCHEME_BFV = 0, SCHEME_CKKS = 1 };

struct Modulus {
    u64 q = 0;
    u64 psi = 0;                  // primitive 2N-th root (Lattigo: g^((q-1)/2N), g smallest generator)
    std::vector<u64> psi_br;      // psi^bitrev(j)
    std::vector<u64> ipsi_br;     // psi^-bitrev(j)
    u64 n_inv = 0;
};

inline Modulus make_modulus(u64 q, int logN) {
    Modulus m;
    u64 N = 1ull << logN;
    if ((q - 1) % (2 * N) != 0 || !is_prime(q)) throw std::runtime_error("modulus is not an NTT prime");
    m.q = q;
    u64 g = primitive_root(q);
    m.psi = powmod(g, (q - 1) / (2 * N), q);
    u64 ipsi = invmod(m.psi, q);
    m.psi_br.resize(N);
    m.ipsi_br.resize(N);
    u64 pw = 1, ipw = 1;
    for (u64 j = 0; j < N; j++) {
        u64 r = bitrev(j, logN);
        m.psi_br[r] = pw;
        m.ipsi_br[r] = ipw;
        pw = mulmod(pw, m.psi, q);
        ipw = mulmod(ipw, ipsi, q);
    }
    m.n_inv = invmod(N % q, q);
    return m;
}

// Lattigo ring.NTT: Cooley-Tukey, natural in, bit-reversed out.
inline void ntt_forward(const Modulus& m, int N, u64* a) {
    u64 q = m.q;
    int t = N;
    for (int mm = 1; mm < N; mm <<= 1) {
        t >>= 1;
        for (int i = 0; i < mm; i++) {
            int j1 = 2 * i * t;
            u64 W = m.psi_br[mm + i];
            for (int j = j1; j < j1 + t; j++) {
                u64 U = a[j];
                u64 V = mulmod(a[j + t], W, q);
                a[j] = addmod(U, V, q);
                a[j + t] = submod(U, V, q);
            }
        }
    }
}

// Lattigo ring.InvNTT: Gentleman-Sande, bit-reversed in, natural out, N^-1 folded.
inline void ntt_inverse(const Modulus& m, int N, u64* a) {
    u64 q = m.q;
    int t = 1;
    for (int mm = N; mm > 1; mm >>= 1) {
        int h = mm >> 1;
        int j1 = 0;
        for (int i = 0; i < h; i++) {
            u64 W = m.ipsi_br[h + i];
            for (int j = j1; j < j1 + t; j++) {
                u64 U = a[j];
                u64 V = a[j + t];
                a[j] = addmod(U, V, q);
                a[j + t] = mulmod(submod(U, V, q), W, q);
            }
            j1 += 2 * t;
        }
        t <<= 1;
    }
    for (int j = 0; j < N; j++) a[j] = mulmod(a[j], m.n_inv, q);
}

struct Context {
    int scheme = SCHEME_CKKS;
    int logN = 0, N = 0;
    int nq = 0, np = 0, nqm = 0;
    u64 t = 0;
    std::vector<Modulus> mods;        // Q..., P..., QMul...
    std::vector<int> level_qmul;      // BFV: levelQMul[level]
    const Modulus& Q(int i) const { return mods[i]; }
    const Modulus& P(int j) const { return mods[nq + j]; }
    const Modulus& QM(int j) const { return mods[nq + np + j]; }
    int idxQ(int i) const { return i; }
    int idxP(int j) const { return nq + j; }
    int idxQM(int j) const { return nq + np + j; }
};

inline std::unique_ptr<Context> make_context(int scheme, int logN, const u64* q, int nq, const u64* p, int np, u64 t) {
    auto c = std::make_unique<Context>();
    c->scheme = scheme;
    c->logN = logN;
    c->N = 1 << logN;
    c->nq = nq;
    c->np = np;
    c->t = t;
    for (int i = 0; i < nq; i++) c->mods.push_back(make_modulus(q[i], logN));
    for (int j = 0; j < np; j++) c->mods.push_back(make_modulus(p[j], logN));
    if (scheme == SCHEME_BFV) {
        // bfv.NewParameters: nbQiMul = ceil((bitlen(Q)+logN)/61); QMul = GenerateNTTPrimesP(61, 2N, nbQiMul)
        BigUInt Qfull(1);
        c->level_qmul.resize(nq);
        for (int i = 0; i < nq; i++) {
            Qfull.mul(q[i]);
            // bfv.NewEvaluator: levelQMul[i] = ceil((bitlen(Q_i)+logN)/61) - 1
            c->level_qmul[i] = (Qfull.bitlen() + logN + 60) / 61 - 1;
        }
        int nqm = (Qfull.bitlen() + logN + 60) / 61;
        std::vector<u64> qm = generate_ntt_primes_p(61, 2ull << logN, nqm);
        for (u64 v : qm) c->mods.push_back(make_modulus(v, logN));
        c->nqm = nqm;
    }
    return c;
}

// ---------------------------------------------------------------------------------------------
// Basis extension (Lattigo ring/basis_extension.go modUpExact / reconstructRNS / multSum):
//   y_i = x_i * (S/s_i)^-1 mod s_i ;  v = uint64(sum_i float64(y_i)/float64(s_i))  (limb order)
//   out_t = sum_i y_i * (S/s_i mod p_t) - v * (S mod p_t)   (mod p_t)
// src[i] / dst[j] point to N-element limbs; sidx/tidx are indices into ctx.mods.
// ---------------------------------------------------------------------------------------------
inline void basis_extend(const Context& ctx, const int* sidx, int ns, const u64* const* src, const int* tidx, int nt,
                         u64* const* dst) {
    int N = ctx.N;
    std::vector<u64> qhat_inv(ns);
    std::vector<std::vector<u64>> qhat_mod(nt, std::vector<u64>(ns));
    std::vector<u64> S_mod(nt);
    for (int i = 0; i < ns; i++) {
        u64 si = ctx.mods[sidx[i]].q;
        u64 prod = 1;
        for (int k = 0; k < ns; k++)
            if (k != i) prod = mulmod(prod, ctx.mods[sidx[k]].q % si, si);
        qhat_inv[i] = invmod(prod, si);
    }
    for (int j = 0; j < nt; j++) {
        u64 pj = ctx.mods[tidx[j]].q;
        u64 all = 1;
        for (int k = 0; k < ns; k++) all = mulmod(all, ctx.mods[sidx[k]].q % pj, pj);
        S_mod[j] = all;
        for (int i = 0; i < ns; i++) {
            u64 prod = 1;
            for (int k = 0; k < ns; k++)
                if (k != i) prod = mulmod(prod, ctx.mods[sidx[k]].q % pj, pj);
            qhat_mod[j][i] = prod;
        }
    }
    std::vector<u64> y(ns);
    for (int x = 0; x < N; x++) {
        double vf = 0.0;
        for (int i = 0; i < ns; i++) {
            u64 si = ctx.mods[sidx[i]].q;
            y[i] = mulmod(src[i][x], qhat_inv[i], si);
            vf += (double)y[i] / (double)si;
        }
        u64 v = (u64)vf;
        for (int j = 0; j < nt; j++) {
            u64 pj = ctx.mods[tidx[j]].q;
            u64 acc = 0;
            for (int i = 0; i < ns; i++) acc = addmod(acc, mulmod(y[i] % pj, qhat_mod[j][i], pj), pj);
            acc = submod(acc, mulmod(v % pj, S_mod[j], pj), pj);
            dst[j][x] = acc;
        }
    }
}

// rlwe DecomposeAndSplit, single-modulus digit: centred lift of x mod s into each target modulus.
inline void centered_lift(const Context& ctx, int sidx, const u64* src, const int* tidx, int nt, u64* const* dst) {
    int N = ctx.N;
    u64 s = ctx.mods[sidx].q;
    for (int x = 0; x < N; x++) {
        u64 c = src[x];
        bool neg = c >= (s >> 1);
        if (neg) c = s - c;
        for (int j = 0; j < nt; j++) {
            u64 pj = ctx.mods[tidx[j]].q;
            u64 r = c % pj;
            dst[j][x] = neg ? negmod(r, pj) : r;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Hybrid key-switch (rlwe.KeySwitcher.SwitchKeysInPlace).
//   c_ntt : [l][N] NTT-domain input polynomial, l = level+1
//   evk   : [beta_key][2][key_level+1+k][N], NTT domain, Montgomery removed (reference ABI form)
//   p0,p1 : [l][N] outputs, NTT domain
// ---------------------------------------------------------------------------------------------
inline void keyswitch_ntt(const Context& ctx, int level, const u64* c_ntt, const u64* evk, int key_level, u64* p0,
                          u64* p1) {
    const int N = ctx.N, l = level + 1, k = ctx.np, alpha = k;
    const int beta = (l + alpha - 1) / alpha;
    const int klimbs = key_level + 1 + k;
    std::vector<u64> c_coeff((size_t)l * N);
    std::memcpy(c_coeff.data(), c_ntt, sizeof(u64) * l * N);
    for (int i = 0; i < l; i++) ntt_inverse(ctx.Q(i), N, &c_coeff[(size_t)i * N]);

    std::vector<u64> acc0((size_t)(l + k) * N, 0), acc1((size_t)(l + k) * N, 0);
    std::vector<u64> d((size_t)(l + k) * N);
    for (int dg = 0; dg < beta; dg++) {
        int st = dg * alpha, ed = std::min(st + alpha, l);
        std::vector<int> sidx, tidx;
        std::vector<const u64*> src;
        std::vector<u64*> dst;
        for (int i = st; i < ed; i++) { sidx.push_back(ctx.idxQ(i)); src.push_back(&c_coeff[(size_t)i * N]); }
        for (int i = 0; i < l; i++)
            if (i < st || i >= ed) { tidx.push_back(ctx.idxQ(i)); dst.push_back(&d[(size_t)i * N]); }
        for (int j = 0; j < k; j++) { tidx.push_back(ctx.idxP(j)); dst.push_back(&d[(size_t)(l + j) * N]); }
        if (ed - st == 1)
            centered_lift(ctx, sidx[0], src[0], tidx.data(), (int)tidx.size(), dst.data());
        else
            basis_extend(ctx, sidx.data(), (int)sidx.size(), src.data(), tidx.data(), (int)tidx.size(), dst.data());
        for (size_t j = 0; j < tidx.size(); j++) ntt_forward(ctx.mods[tidx[j]], N, dst[j]);
        for (int i = st; i < ed; i++) std::memcpy(&d[(size_t)i * N], &c_ntt[(size_t)i * N], sizeof(u64) * N);
        // acc_b[t] += d[t] * evk[dg][b][t]
        for (int tl = 0; tl < l + k; tl++) {
            int key_limb = tl < l ? tl : key_level + 1 + (tl - l);
            u64 q = tl < l ? ctx.Q(tl).q : ctx.P(tl - l).q;
            const u64* e0 = evk + ((size_t)(dg * 2 + 0) * klimbs + key_limb) * N;
            const u64* e1 = evk + ((size_t)(dg * 2 + 1) * klimbs + key_limb) * N;
            const u64* dd = &d[(size_t)tl * N];
            u64* a0 = &acc0[(size_t)tl * N];
            u64* a1 = &acc1[(size_t)tl * N];
            for (int x = 0; x < N; x++) {
                a0[x] = addmod(a0[x], mulmod(dd[x], e0[x], q), q);
                a1[x] = addmod(a1[x], mulmod(dd[x], e1[x], q), q);
            }
        }
    }
    // ModDownQPtoQNTT on both accumulators
    for (int b = 0; b < 2; b++) {
        std::vector<u64>& acc = b ? acc1 : acc0;
        u64* out = b ? p1 : p0;
        std::vector<int> sidx, tidx;
        std::vector<const u64*> src;
        std::vector<u64*> dst;
        std::vector<u64> ext((size_t)l * N);
        for (int j = 0; j < k; j++) {
            ntt_inverse(ctx.P(j), N, &acc[(size_t)(l + j) * N]);
            sidx.push_back(ctx.idxP(j));
            src.push_back(&acc[(size_t)(l + j) * N]);
        }
        for (int i = 0; i < l; i++) { tidx.push_back(ctx.idxQ(i)); dst.push_back(&ext[(size_t)i * N]); }
        basis_extend(ctx, sidx.data(), k, src.data(), tidx.data(), l, dst.data());
        for (int i = 0; i < l; i++) {
            u64 q = ctx.Q(i).q;
            ntt_forward(ctx.Q(i), N, dst[i]);
            u64 pinv = 1;
            for (int j = 0; j < k; j++) pinv = mulmod(pinv, ctx.P(j).q % q, q);
            pinv = invmod(pinv, q);
            for (int x = 0; x < N; x++)
                out[(size_t)i * N + x] = mulmod(submod(acc[(size_t)i * N + x], dst[i][x], q), pinv, q);
        }
    }
}

// Galois permutation tables.
inline std::vector<uint32_t> permute_ntt_index(int logN, u64 gal) {  // ring.PermuteNTTIndex
    u64 N = 1ull << logN, mask = 2 * N - 1;
    std::vector<uint32_t> idx(N);
    for (u64 i = 0; i < N; i++) {
        u64 t1 = 2 * bitrev(i, logN) + 1;
        u64 t2 = ((gal * t1) & mask) - 1;
        idx[i] = (uint32_t)bitrev(t2 >> 1, logN);
    }
    return idx;
}
inline void permute_ntt(int N, const std::vector<uint32_t>& idx, const u64* in, u64* out) {
    for (int i = 0; i < N; i++) out[i] = in[idx[i]];
}
// ring.Permute (coefficient domain): X^i -> X^(i*gal)
inline void permute_coeff(int logN, u64 gal, u64 q, const u64* in, u64* out) {
    u64 N = 1ull << logN, mask = N - 1;
    for (u64 i = 0; i < N; i++) {
        u64 index = i * gal;
        u64 sign = (index >> logN) & 1;
        out[index & mask] = sign ? negmod(in[i], q) : in[i];
    }
}

// ---------------------------------------------------------------------------------------------
// Limb-wise helpers over [polys][l][N] blocks.
// ---------------------------------------------------------------------------------------------
inline void ct_add(const Context& ctx, int level, int npoly, const u64* a, const u64* b, u64* o) {
    int N = ctx.N, l = level + 1;
    for (int p = 0; p < npoly; p++)
        for (int i = 0; i < l; i++) {
            u64 q = ctx.Q(i).q;
            size_t off = ((size_t)p * l + i) * N;
            for (int x = 0; x < N; x++) o[off + x] = addmod(a[off + x], b[off + x], q);
        }
}
inline void ct_sub(const Context& ctx, int level, int npoly, const u64* a, const u64* b, u64* o) {
    int N = ctx.N, l = level + 1;
    for (int p = 0; p < npoly; p++)
        for (int i = 0; i < l; i++) {
            u64 q = ctx.Q(i).q;
            size_t off = ((size_t)p * l + i) * N;
            for (int x = 0; x < N; x++) o[off + x] = submod(a[off + x], b[off + x], q);
        }
}
inline void ct_neg(const Context& ctx, int level, int npoly, const u64* a, u64* o) {
    int N = ctx.N, l = level + 1;
    for (int p = 0; p < npoly; p++)
        for (int i = 0; i < l; i++) {
            u64 q = ctx.Q(i).q;
            size_t off = ((size_t)p * l + i) * N;
            for (int x = 0; x < N; x++) o[off + x] = negmod(a[off + x], q);
        }
}

// NTT-domain degree-1 x degree-1 tensor (ckks evaluator mulRelin without relin): [2][l][N]x2 -> [3][l][N]
inline void tensor_ntt(const Context& ctx, const int* midx, int l, const u64* a, const u64* b, u64* o) {
    int N = ctx.N;
    for (int i = 0; i < l; i++) {
        u64 q = ctx.mods[midx[i]].q;
        const u64 *a0 = a + (size_t)i * N, *a1 = a + (size_t)(l + i) * N;
        const u64 *b0 = b + (size_t)i * N, *b1 = b + (size_t)(l + i) * N;
        u64 *o0 = o + (size_t)i * N, *o1 = o + (size_t)(l + i) * N, *o2 = o + (size_t)(2 * l + i) * N;
        for (int x = 0; x < N; x++) {
            u64 c0 = mulmod(a0[x], b0[x], q);
            u64 c1 = addmod(mulmod(a0[x], b1[x], q), mulmod(a1[x], b0[x], q), q);
            u64 c2 = mulmod(a1[x], b1[x], q);
            o0[x] = c0; o1[x] = c1; o2[x] = c2;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// CKKS
// ---------------------------------------------------------------------------------------------
inline void ckks_mult(const Context& ctx, int level, const u64* a, const u64* b, u64* out3) {
    std::vector<int> midx(level + 1);
    for (int i = 0; i <= level; i++) midx[i] = i;
    tensor_ntt(ctx, midx.data(), level + 1, a, b, out3);
}

// ckks Relinearize: (c0,c1) + KS(c2)
inline void ckks_relin(const Context& ctx, int level, const u64* ct3, const u64* rlk, int key_level, u64* out2) {
    int N = ctx.N, l = level + 1;
    std::vector<u64> p0((size_t)l * N), p1((size_t)l * N);
    keyswitch_ntt(ctx, level, ct3 + (size_t)2 * l * N, rlk, key_level, p0.data(), p1.data());
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        for (int x = 0; x < N; x++) {
            out2[(size_t)i * N + x] = addmod(ct3[(size_t)i * N + x], p0[(size_t)i * N + x], q);
            out2[(size_t)(l + i) * N + x] = addmod(ct3[(size_t)(l + i) * N + x], p1[(size_t)i * N + x], q);
        }
    }
}

// ring.DivRoundByLastModulusNTT on each polynomial: [np][l][N] -> [np][l-1][N]
inline void ckks_rescale(const Context& ctx, int level, int npoly, const u64* in, u64* out) {
    int N = ctx.N, l = level + 1;
    u64 ql = ctx.Q(level).q, half = (ql - 1) >> 1;
    std::vector<u64> last(N), tmp(N);
    for (int p = 0; p < npoly; p++) {
        std::memcpy(last.data(), in + ((size_t)p * l + level) * N, sizeof(u64) * N);
        ntt_inverse(ctx.Q(level), N, last.data());
        for (int x = 0; x < N; x++) last[x] = addmod(last[x], half, ql);
        for (int i = 0; i < level; i++) {
            u64 q = ctx.Q(i).q;
            u64 half_q = half % q;
            u64 qlinv = invmod(ql % q, q);
            for (int x = 0; x < N; x++) tmp[x] = submod(last[x] % q, half_q, q);
            ntt_forward(ctx.Q(i), N, tmp.data());
            const u64* src = in + ((size_t)p * l + i) * N;
            u64* dst = out + ((size_t)p * (l - 1) + i) * N;
            for (int x = 0; x < N; x++) dst[x] = mulmod(submod(src[x], tmp[x], q), qlinv, q);
        }
    }
}

inline void drop_level(const Context& ctx, int level, int npoly, const u64* in, u64* out) {
    int N = ctx.N, l = level + 1;
    for (int p = 0; p < npoly; p++)
        std::memcpy(out + (size_t)p * (l - 1) * N, in + (size_t)p * l * N, sizeof(u64) * (l - 1) * N);
}

// ckks evaluator.permuteNTT: key-switch c1 under glk[gal], add c0, then permute both.
inline void ckks_rotate(const Context& ctx, int level, const u64* ct, const u64* glk, int key_level, u64 gal,
                        u64* out) {
    int N = ctx.N, l = level + 1;
    std::vector<u64> p0((size_t)l * N), p1((size_t)l * N);
    keyswitch_ntt(ctx, level, ct + (size_t)l * N, glk, key_level, p0.data(), p1.data());
    std::vector<uint32_t> idx = permute_ntt_index(ctx.logN, gal);
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        for (int x = 0; x < N; x++) p0[(size_t)i * N + x] = addmod(p0[(size_t)i * N + x], ct[(size_t)i * N + x], q);
        permute_ntt(N, idx, &p0[(size_t)i * N], out + (size_t)i * N);
        permute_ntt(N, idx, &p1[(size_t)i * N], out + (size_t)(l + i) * N);
    }
}

// ---------------------------------------------------------------------------------------------
// BFV (coefficient-domain ciphertexts)
// ---------------------------------------------------------------------------------------------
// bfv evaluator.tensorAndRescale: ModUp Q->QMul, NTT, tensor in both bases, INTT, quantize.
inline void bfv_mult(const Context& ctx, int level, const u64* a, const u64* b, u64* out3) {
    const int N = ctx.N, l = level + 1, lm = ctx.level_qmul[level] + 1, L = l + lm;
    std::vector<int> midx(L), qidx(l), midx_m(lm);
    for (int i = 0; i < l; i++) midx[i] = qidx[i] = ctx.idxQ(i);
    for (int j = 0; j < lm; j++) midx[l + j] = midx_m[j] = ctx.idxQM(j);
    auto extend_and_ntt = [&](const u64* ct, std::vector<u64>& e) {
        e.assign((size_t)2 * L * N, 0);
        for (int p = 0; p < 2; p++) {
            std::vector<const u64*> src;
            std::vector<u64*> dst;
            for (int i = 0; i < l; i++) {
                std::memcpy(&e[((size_t)p * L + i) * N], ct + ((size_t)p * l + i) * N, sizeof(u64) * N);
                src.push_back(ct + ((size_t)p * l + i) * N);
            }
            for (int j = 0; j < lm; j++) dst.push_back(&e[((size_t)p * L + l + j) * N]);
            basis_extend(ctx, qidx.data(), l, src.data(), midx_m.data(), lm, dst.data());
            for (int i = 0; i < L; i++) ntt_forward(ctx.mods[midx[i]], N, &e[((size_t)p * L + i) * N]);
        }
    };
    std::vector<u64> ea, eb, prod((size_t)3 * L * N);
    extend_and_ntt(a, ea);
    extend_and_ntt(b, eb);
    tensor_ntt(ctx, midx.data(), L, ea.data(), eb.data(), prod.data());
    // pHalf = QMul_level / 2
    BigUInt QM(1);
    for (int j = 0; j < lm; j++) QM.mul(ctx.QM(j).q);
    BigUInt half = QM;
    half.shr1();
    std::vector<u64> buf((size_t)lm * N), ext((size_t)lm * N);
    for (int p = 0; p < 3; p++) {
        u64* pp = &prod[(size_t)p * L * N];
        for (int i = 0; i < L; i++) ntt_inverse(ctx.mods[midx[i]], N, pp + (size_t)i * N);
        // ModDownQPtoP: (xP - ext([x]_Q)) * Q^-1 mod each QMul prime
        std::vector<const u64*> src;
        std::vector<u64*> dst;
        for (int i = 0; i < l; i++) src.push_back(pp + (size_t)i * N);
        for (int j = 0; j < lm; j++) dst.push_back(&ext[(size_t)j * N]);
        basis_extend(ctx, qidx.data(), l, src.data(), midx_m.data(), lm, dst.data());
        for (int j = 0; j < lm; j++) {
            u64 m = ctx.QM(j).q;
            u64 qinv = 1;
            for (int i = 0; i < l; i++) qinv = mulmod(qinv, ctx.Q(i).q % m, m);
            qinv = invmod(qinv, m);
            u64 hm = half.mod(m);
            const u64* xp = pp + (size_t)(l + j) * N;
            for (int x = 0; x < N; x++) {
                u64 v = mulmod(submod(xp[x], ext[(size_t)j * N + x], m), qinv, m);
                buf[(size_t)j * N + x] = addmod(v, hm, m);  // AddScalarBigint(pHalf)
            }
        }
        // ModUpPtoQ (QMul -> Q), subtract pHalf, multiply by t
        std::vector<const u64*> src2;
        std::vector<u64*> dst2;
        for (int j = 0; j < lm; j++) src2.push_back(&buf[(size_t)j * N]);
        for (int i = 0; i < l; i++) dst2.push_back(out3 + ((size_t)p * l + i) * N);
        basis_extend(ctx, midx_m.data(), lm, src2.data(), qidx.data(), l, dst2.data());
        for (int i = 0; i < l; i++) {
            u64 q = ctx.Q(i).q;
            u64 hq = half.mod(q);
            u64 tq = ctx.t % q;
            u64* o = out3 + ((size_t)p * l + i) * N;
            for (int x = 0; x < N; x++) o[x] = mulmod(submod(o[x], hq, q), tq, q);
        }
    }
}

inline void bfv_keyswitch_coeff(const Context& ctx, int level, const u64* c_coeff, const u64* evk, int key_level,
                                u64* p0, u64* p1) {
    int N = ctx.N, l = level + 1;
    std::vector<u64> c_ntt(c_coeff, c_coeff + (size_t)l * N);
    for (int i = 0; i < l; i++) ntt_forward(ctx.Q(i), N, &c_ntt[(size_t)i * N]);
    keyswitch_ntt(ctx, level, c_ntt.data(), evk, key_level, p0, p1);
    for (int i = 0; i < l; i++) {
        ntt_inverse(ctx.Q(i), N, p0 + (size_t)i * N);
        ntt_inverse(ctx.Q(i), N, p1 + (size_t)i * N);
    }
}

inline void bfv_relin(const Context& ctx, int level, const u64* ct3, const u64* rlk, int key_level, u64* out2) {
    int N = ctx.N, l = level + 1;
    std::vector<u64> p0((size_t)l * N), p1((size_t)l * N);
    bfv_keyswitch_coeff(ctx, level, ct3 + (size_t)2 * l * N, rlk, key_level, p0.data(), p1.data());
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        for (int x = 0; x < N; x++) {
            out2[(size_t)i * N + x] = addmod(ct3[(size_t)i * N + x], p0[(size_t)i * N + x], q);
            out2[(size_t)(l + i) * N + x] = addmod(ct3[(size_t)(l + i) * N + x], p1[(size_t)i * N + x], q);
        }
    }
}

// ring.DivRoundByLastModulus (coefficient domain) — BFV "rescale" = modulus switch by one level.
inline void bfv_rescale(const Context& ctx, int level, int npoly, const u64* in, u64* out) {
    int N = ctx.N, l = level + 1;
    u64 ql = ctx.Q(level).q, half = (ql - 1) >> 1;
    for (int p = 0; p < npoly; p++) {
        const u64* last = in + ((size_t)p * l + level) * N;
        for (int i = 0; i < level; i++) {
            u64 q = ctx.Q(i).q;
            u64 half_q = half % q;
            u64 qlinv = invmod(ql % q, q);
            const u64* src = in + ((size_t)p * l + i) * N;
            u64* dst = out + ((size_t)p * (l - 1) + i) * N;
            for (int x = 0; x < N; x++) {
                u64 sh = addmod(last[x], half, ql);
                u64 lift = submod(sh % q, half_q, q);
                dst[x] = mulmod(submod(src[x], lift, q), qlinv, q);
            }
        }
    }
}

// bfv evaluator rotate: key-switch c1, add c0, then ring.Permute both.
inline void bfv_rotate(const Context& ctx, int level, const u64* ct, const u64* glk, int key_level, u64 gal, u64* out) {
    int N = ctx.N, l = level + 1;
    std::vector<u64> p0((size_t)l * N), p1((size_t)l * N);
    bfv_keyswitch_coeff(ctx, level, ct + (size_t)l * N, glk, key_level, p0.data(), p1.data());
    for (int i = 0; i < l; i++) {
        u64 q = ctx.Q(i).q;
        for (int x = 0; x < N; x++) p0[(size_t)i * N + x] = addmod(p0[(size_t)i * N + x], ct[(size_t)i * N + x], q);
        permute_coeff(ctx.logN, gal, q, &p0[(size_t)i * N], out + (size_t)i * N);
        permute_coeff(ctx.logN, gal, q, &p1[(size_t)i * N], out + (size_t)(l + i) * N);
    }
}

}  // namespace orc
