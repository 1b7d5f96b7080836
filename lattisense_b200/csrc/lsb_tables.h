// Host-only builders of the per-modulus NTT tables and CRT conversion tables (pure C++; shared by
// the CUDA context and by tests/emul).
#pragma once
#include <cstring>
#include <vector>
#include "lsb_kernels.h"
#include "lsb_host_math.h"

namespace lsb {

// fwd/inv: N entries each.  psi = g^((q-1)/2N), g the smallest generator of Z_q^* (Lattigo ring.NTT
// parameters); tables hold psi^bitrev(j) / psi^-bitrev(j) with Shoup companions.
inline TwD make_twd(u64 w, u64 q) { return TwD{(double)w, (double)w / (double)q}; }

inline void build_ntt_tables(u64 qq, int logN, Tw* f, Tw* iv, ModConst* mc, TwD* fd = nullptr, TwD* ivd = nullptr) {
    using namespace host;
    const u64 N = 1ull << logN;
    LSB_CHECK(qq < (1ull << 62), "lsb: modulus must be below 2^62");
    LSB_CHECK((qq - 1) % (2 * N) == 0 && miller_rabin(qq), "lsb: modulus is not an NTT-friendly prime");
    const u64 g = smallest_generator(qq);
    const u64 psi = powmod(g, (qq - 1) / (2 * N), qq);
    const u64 ipsi = invmod(psi, qq);
    u64 pw = 1, ipw = 1;
    for (u64 j = 0; j < N; j++) {
        const u64 r = bitrev(j, logN);
        f[r] = make_tw(pw, qq);
        iv[r] = make_tw(ipw, qq);
        pw = mulmod(pw, psi, qq);
        ipw = mulmod(ipw, ipsi, qq);
    }
    const u64 ninv = invmod(N % qq, qq);
    mc->q = qq;
    mc->qinv64 = ~0ull / qq;
    mc->bq = make_barrett(qq);
    mc->ninv = make_tw(ninv, qq);
    mc->w_ninv = make_tw(mulmod(iv[1].w, ninv, qq), qq);
    mc->f64_ok = f64_eligible(qq, logN) ? 1 : 0;
    mc->f64_recentre = f64_needs_recentre(qq, logN) ? 1 : 0;
    mc->qd = (double)qq;
    mc->qinvd = 1.0 / (double)qq;
    mc->ninv_d = make_twd(mc->ninv.w, qq);
    mc->w_ninv_d = make_twd(mc->w_ninv.w, qq);
    if (fd && ivd)
        for (u64 j = 0; j < N; j++) {
            fd[j] = mc->f64_ok ? make_twd(f[j].w, qq) : TwD{0.0, 0.0};
            ivd[j] = mc->f64_ok ? make_twd(iv[j].w, qq) : TwD{0.0, 0.0};
        }
}

// Re-order the stages s1+1 .. logN-1 (s1 = logN - lrl, lrl = radix of the cluster kernel's contiguous last round)
// of one modulus' bit-reversed twiddle table into the tail-SoA layout read through tw_index<S1> (lsb_ntt_arith.h):
// entry (tl << r) | kg of stage s1 + r moves to kg * 2^s1 + tl.
template <class T>
inline void soa_tail(T* table, int logN, int lrl) {
    const int s1 = logN - lrl;
    std::vector<T> seg;
    for (int r = 1; r < lrl; r++) {
        const size_t base = (size_t)1 << (s1 + r), groups = (size_t)1 << s1;
        seg.assign(table + base, table + 2 * base);
        for (size_t tl = 0; tl < groups; tl++)
            for (size_t kg = 0; kg < ((size_t)1 << r); kg++) table[base + kg * groups + tl] = seg[(tl << r) | kg];
    }
}

// Mixed-pipe CRT of a job that cannot run on the FP64 pipe as a whole (a source or a target modulus >= 2^47): choose which
// of the eligible targets (modulus < 2^47, sum of the job's terms below 2^51) go to the FP64 pipe so that the two pipes
// finish together.  Per-thread cost model from the micro-benchmarks (profiles/r01_int_pipe_microbench.txt): a 64x64->128
// multiply-accumulate 0.22 clk/SM-lane, a reduction 0.35; an FP64-pipe instruction 1/57.8, 6 per product.
inline void bex_choose_mixed(BexJob& j, const std::vector<u64>& moduli) {
    using namespace host;
    j.n_f64 = 0;
    j.src_split = 0;
    std::memset(j.t_f64, 0, sizeof(j.t_f64));
    static const double bias = [] {  // LSB_BEX_MIXED=0: integer pipe only; LSB_BEX_F64_BIAS=x: scale the FP64 cost
        const char* off = getenv("LSB_BEX_MIXED");
        if (off && atoi(off) == 0) return -1.0;
        const char* b = getenv("LSB_BEX_F64_BIAS");
        return b ? atof(b) : 1.0;
    }();
    if (bias <= 0.0 || j.f64_ok || j.ns < 1) return;
    for (int i = 0; i < j.ns; i++)
        if (moduli[j.src_mod[i]] >= (1ull << 51)) j.src_split = 1;
    const int terms = j.ns * (j.src_split ? 2 : 1);
    std::vector<int> elig;
    for (int t = 0; t < j.nt; t++) {
        const u64 p = moduli[j.dst_mod[t]];
        // |sum| <= 0.75 p * terms, then - vS (< p), then the mode's arithmetic: keep everything below 2^51
        if (p < kElemF64Limit && (u128)p * (u128)(terms + 2) < ((u128)1 << 51)) elig.push_back(t);
    }
    const double cf = bias * (terms * 6.0 + 10.0) / 57.8, ci = j.ns * 0.22 + 0.35;
    int best = 0;
    double best_cost = j.nt * ci;
    for (int nf = 1; nf <= (int)elig.size(); nf++) {
        const double c = std::max(nf * cf, (j.nt - nf) * ci);
        if (c < best_cost) {
            best_cost = c;
            best = nf;
        }
    }
    // spread the FP64 targets over the job so that consecutive targets alternate pipes
    for (int n = 0; n < best; n++) {
        const int t = elig[(size_t)n * elig.size() / best];
        j.t_f64[t] = 1;
    }
    j.n_f64 = best;
    for (int t = 0; t < j.nt; t++) {
        if (!j.t_f64[t]) continue;
        const u64 p = moduli[j.dst_mod[t]];
        for (int i = 0; i < j.ns; i++) {
            j.qhat_mod_d[t][i] = make_twd(j.qhat_mod[t][i], p);
            j.qhat_mod_hi_d[t][i] = make_twd(mulmod(j.qhat_mod[t][i], (1ull << 32) % p, p), p);
        }
        j.mulc_d[t] = make_twd(j.mulc[t].w, p);
    }
}

// FP64 copies of a job's constants (call again after changing mulc): valid when every modulus is below 2^47
inline void bex_fill_f64(BexJob& j, const std::vector<u64>& moduli) {
    j.f64_ok = 1;
    for (int i = 0; i < j.ns; i++) j.f64_ok &= moduli[j.src_mod[i]] < kElemF64Limit ? 1 : 0;
    for (int t = 0; t < j.nt; t++) j.f64_ok &= moduli[j.dst_mod[t]] < kElemF64Limit ? 1 : 0;
    if (!j.f64_ok) {
        bex_choose_mixed(j, moduli);  // the job stays on the integer path; some of its targets may still use the FP64 pipe
        return;
    }
    j.n_f64 = 0;
    for (int i = 0; i < j.ns; i++) j.qhat_inv_d[i] = make_twd(j.qhat_inv[i].w, moduli[j.src_mod[i]]);
    for (int t = 0; t < j.nt; t++) {
        const u64 p = moduli[j.dst_mod[t]];
        for (int i = 0; i < j.ns; i++) j.qhat_mod_d[t][i] = make_twd(j.qhat_mod[t][i], p);
        j.mulc_d[t] = make_twd(j.mulc[t].w, p);
    }
}

// CRT tables for extending residues from the moduli `src` to each modulus of `dst` (indices into `moduli`).
inline BexJob make_bex_tables(const std::vector<u64>& moduli, const std::vector<int>& src, const std::vector<int>& dst) {
    using namespace host;
    BexJob j;
    std::memset(&j, 0, sizeof(j));
    LSB_CHECK((int)src.size() <= kBexMaxSrc && (int)dst.size() <= kBexMaxDst, "lsb: basis extension too wide");
    j.ns = (int)src.size();
    j.nt = (int)dst.size();
    j.mode = BEX_PLAIN;
    {
        u64 smax = 0;
        for (int i = 0; i < j.ns; i++) smax = std::max(smax, moduli[src[i]]);
        j.wide_ok = ((u128)smax * (u128)j.ns < ((u128)1 << 63)) ? 1 : 0;
    }
    for (int i = 0; i < j.ns; i++) {
        const u64 si = moduli[src[i]];
        j.src_mod[i] = (uint8_t)src[i];
        j.s_f64[i] = (double)si;
        j.rs_f64[i] = 1.0 / (double)si;
        u64 prod = 1;
        for (int k = 0; k < j.ns; k++)
            if (k != i) prod = mulmod(prod, moduli[src[k]] % si, si);
        j.qhat_inv[i] = make_tw(invmod(prod, si), si);
    }
    for (int tt = 0; tt < j.nt; tt++) {
        const u64 pj = moduli[dst[tt]];
        j.dst_mod[tt] = (uint8_t)dst[tt];
        u64 all = 1;
        for (int k = 0; k < j.ns; k++) all = mulmod(all, moduli[src[k]] % pj, pj);
        for (int v = 0; v <= j.ns; v++) j.vS[tt][v] = mulmod((u64)v % pj, all, pj);
        for (int i = 0; i < j.ns; i++) {
            u64 prod = 1;
            for (int k = 0; k < j.ns; k++)
                if (k != i) prod = mulmod(prod, moduli[src[k]] % pj, pj);
            j.qhat_mod[tt][i] = prod;
        }
    }
    bex_fill_f64(j, moduli);
    return j;
}

}  // namespace lsb
