// TEST INFRASTRUCTURE: runs the LSB_HD round / per-coefficient code of lattisense_b200/csrc on the
// CPU (thread loops in place of CTAs) so that indexing, twiddle selection, lazy ranges and the
// CRT tables can be checked against the oracle without a GPU.  Never used by the product path.
#include <vector>
#include <cstring>
#include "../../lattisense_b200/csrc/lsb_ntt_core.h"
#include "../../lattisense_b200/csrc/lsb_tables.h"
#include "../../lattisense_b200/csrc/lsb_elem_core.h"
#include "../../lattisense_b200/csrc/lsb_ntt_cluster_core.h"
#include "../../lattisense_b200/csrc/lsb_ntt_persist_core.h"

using namespace lsb;

template <bool INV>
static void round_dyn(int lr, int src, int dst, bool fin, bool vec, u64* sm, const u64* gsrc, u64* gdst, TileGeom tg,
                      const Tw* tw, int g, int loggc, int v, u32 twbase, const ModConst& mc) {
#define CASE(LR, S, D, F, V)                                                                                   \
    if (lr == LR && src == S && dst == D && fin == F && vec == V) {                                            \
        if (!INV && mc.q < kLazyLimit)                                                                         \
            fwd_round<LR, S, D, F, V, true>(sm, gsrc, gdst, tg, tw, g, loggc, v, twbase, mc.q, mc.qinv64);     \
        else if (!INV) fwd_round<LR, S, D, F, V, false>(sm, gsrc, gdst, tg, tw, g, loggc, v, twbase, mc.q);    \
        else inv_round<LR, S, D, F, V>(sm, gsrc, gdst, tg, tw, g, loggc, v, twbase, mc.q, mc.ninv, mc.w_ninv); \
        return;                                                                                                \
    }
#define CASES_LR(LR)                                                                         \
    CASE(LR, 0, 0, false, false) CASE(LR, 0, 0, false, true) CASE(LR, 0, 0, true, false)     \
    CASE(LR, 0, 0, true, true) CASE(LR, 1, 0, false, false) CASE(LR, 0, 1, false, false)     \
    CASE(LR, 0, 1, true, false) CASE(LR, 1, 0, true, false) CASE(LR, 1, 1, false, false) CASE(LR, 1, 1, true, false)
    CASES_LR(1) CASES_LR(2) CASES_LR(3) CASES_LR(4)
    throw std::runtime_error("emul: unhandled round configuration");
}

static void single_pass(int logN, const Tw* f, const Tw* iv, const ModConst& mc, u64* data, bool inverse, int NT) {
    const int N = 1 << logN;
    std::vector<u64> sm(N);
    const Plan P = make_plan(logN, 0);
    const TileGeom tg{0, 0};
    if (!inverse) {
        for (int i = 0; i < P.nrounds; i++) {
            const int lr = P.rounds[i].lr, loggc = plan_loggc(P, i);
            for (int tid = 0; tid < NT; tid++)
                for (int g = tid; g < (N >> lr); g += NT)
                    round_dyn<false>(lr, i == 0 ? IO_GLOBAL : IO_SMEM, IO_SMEM, i == P.nrounds - 1, loggc == 0, sm.data(),
                                     data, nullptr, tg, f, g, loggc, P.rounds[i].v, 1u, mc);
        }
        for (int c = 0; c < N / 2; c++) {
            data[2 * c] = sm[swz(2 * c)];
            data[2 * c + 1] = sm[swz(2 * c) + 1];
        }
    } else {
        for (int c = 0; c < N / 2; c++) {
            sm[swz(2 * c)] = data[2 * c];
            sm[swz(2 * c) + 1] = data[2 * c + 1];
        }
        for (int j = 0; j < P.nrounds; j++) {
            const int i = P.nrounds - 1 - j;
            const int lr = P.rounds[i].lr, loggc = plan_loggc(P, i);
            for (int tid = 0; tid < NT; tid++)
                for (int g = tid; g < (N >> lr); g += NT)
                    round_dyn<true>(lr, IO_SMEM, i == 0 ? IO_GLOBAL : IO_SMEM, i == 0, loggc == 0, sm.data(), nullptr,
                                    data, tg, iv, g, loggc, P.rounds[i].v, 1u, mc);
        }
    }
}

// mirrors k_ntt_cols / k_ntt_rows (kRowLS = 8, kColLC = 4, kRowSub = 16)
static void cols_pass(int logN, const Tw* tw, const ModConst& mc, u64* data, bool inverse) {
    const int LA = logN - 8, LC = 4, T = 1 << (LA + LC), NT = T / 16, UNITS = 256 >> LC;
    const Plan P = make_plan(LA, LC);
    const TileGeom tg{LC, 8};
    std::vector<u64> sm(T);
    for (int cb = 0; cb < UNITS; cb++) {
        u64* base = data + (cb << LC);
        for (int j = 0; j < P.nrounds; j++) {
            const int i = inverse ? P.nrounds - 1 - j : j;
            const int lr = P.rounds[i].lr, loggc = plan_loggc(P, i);
            const int src = j == 0 ? IO_GLOBAL : IO_SMEM, dst = j == P.nrounds - 1 ? IO_GLOBAL : IO_SMEM;
            // a round reading global and writing global in place is only safe because reads precede writes per group
            for (int tid = 0; tid < NT; tid++)
                for (int g = tid; g < (T >> lr); g += NT) {
                    if (!inverse) round_dyn<false>(lr, src, dst, false, false, sm.data(), base, base, tg, tw, g, loggc, P.rounds[i].v, 1u, mc);
                    else round_dyn<true>(lr, src, dst, i == 0, false, sm.data(), base, base, tg, tw, g, loggc, P.rounds[i].v, 1u, mc);
                }
        }
    }
}
static void rows_pass(int logN, const Tw* tw, const ModConst& mc, u64* data, bool inverse) {
    const int N = 1 << logN, LA = logN - 8, S = 256, SUB = 16, T = S * SUB, UNITS = N / T;
    const TileGeom tg{0, 0};
    std::vector<u64> sm(T);
    for (int tb = 0; tb < UNITS; tb++) {
        u64* base = data + (size_t)tb * T;
        if (inverse)
            for (int c = 0; c < T / 2; c++) {
                sm[swz(2 * c)] = base[2 * c];
                sm[swz(2 * c) + 1] = base[2 * c + 1];
            }
        for (int phase = 0; phase < 2; phase++)
            for (int tid = 0; tid < 256; tid++) {
                const int st = tid >> 4, g = tid & 15;
                const u32 twbase = (1u << LA) + (u32)(tb * SUB + st);
                u64* smt = sm.data() + st * S;
                if (!inverse) {
                    if (phase == 0) round_dyn<false>(4, IO_GLOBAL, IO_SMEM, false, false, smt, base + st * S, nullptr, tg, tw, g, 4, 0, twbase, mc);
                    else round_dyn<false>(4, IO_SMEM, IO_SMEM, true, true, smt, nullptr, nullptr, tg, tw, g, 0, 4, twbase, mc);
                } else {
                    if (phase == 0) round_dyn<true>(4, IO_SMEM, IO_SMEM, false, true, smt, nullptr, nullptr, tg, tw, g, 0, 4, twbase, mc);
                    else round_dyn<true>(4, IO_SMEM, IO_GLOBAL, false, false, smt, nullptr, base + st * S, tg, tw, g, 4, 0, twbase, mc);
                }
            }
        if (!inverse)
            for (int c = 0; c < T / 2; c++) {
                base[2 * c] = sm[swz(2 * c)];
                base[2 * c + 1] = sm[swz(2 * c) + 1];
            }
    }
}

// mirrors k_ntt_cluster (lsb_ntt_cluster.cu): cross radix-16 round + local rounds per 1/16 block; the CTA
// split is irrelevant on the CPU (blocks are independent), the index / twiddle / arithmetic code is the same
template <class P, int LR, bool INV, bool RAW, bool CANON, int S1>
static void local_all_s(int logN, int i_round, const Plan& PL, u64* sm, const typename P::TW* tw, const ModCtx& m, u64* gio) {
    const int BS = (1 << logN) / 16, GPB = BS >> LR, loggc = plan_loggc(PL, i_round);
    for (int blk = 0; blk < 16; blk++)
        for (int g = 0; g < GPB; g++)
            local_round<P, LR, INV, RAW, CANON, S1>(sm + blk * BS, tw, g, loggc, PL.rounds[i_round].v, 16u + (u32)blk, m,
                                                    gio + blk * BS, gio + blk * BS);
}
// the contiguous last round (RAW for the inverse, CANON for the forward transform) reads the tail-SoA twiddle layout
template <class P, int LR, bool INV, bool RAW, bool CANON>
static void local_all(int logN, int i_round, const Plan& PL, u64* sm, const typename P::TW* tw, const ModCtx& m, u64* gio) {
    if (!(RAW || CANON)) return local_all_s<P, LR, INV, RAW, CANON, -1>(logN, i_round, PL, sm, tw, m, gio);
    switch (logN - LR) {
        case 8: return local_all_s<P, LR, INV, RAW, CANON, 8>(logN, i_round, PL, sm, tw, m, gio);
        case 9: return local_all_s<P, LR, INV, RAW, CANON, 9>(logN, i_round, PL, sm, tw, m, gio);
        case 10: return local_all_s<P, LR, INV, RAW, CANON, 10>(logN, i_round, PL, sm, tw, m, gio);
        case 11: return local_all_s<P, LR, INV, RAW, CANON, 11>(logN, i_round, PL, sm, tw, m, gio);
        case 12: return local_all_s<P, LR, INV, RAW, CANON, 12>(logN, i_round, PL, sm, tw, m, gio);
        case 13: return local_all_s<P, LR, INV, RAW, CANON, 13>(logN, i_round, PL, sm, tw, m, gio);
        default: throw std::runtime_error("emul: unhandled last-round stage");
    }
}
template <class P, bool INV>
static void local_dyn(int logN, int i, const Plan& PL, u64* sm, const typename P::TW* tw, const ModCtx& m, bool raw, bool canon,
                      u64* gio) {
    const int lr = PL.rounds[i].lr;
#define LCASE(LR, RAW, CANON) \
    if (lr == LR && raw == RAW && canon == CANON) return local_all<P, LR, INV, RAW, CANON>(logN, i, PL, sm, tw, m, gio);
    LCASE(3, false, false) LCASE(3, true, false) LCASE(3, false, true) LCASE(4, false, false) LCASE(4, true, false) LCASE(4, false, true)
    throw std::runtime_error("emul: unhandled cluster round");
}
template <class P>
static void cluster_ntt(int logN, const typename P::TW* tw, const ModConst& mc, typename P::TW ninv, typename P::TW w_ninv, u64* data,
                        bool inverse) {
    const int N = 1 << logN, BS = N / 16;
    const Plan PL = make_plan(logN - 4, 0);
    ModCtx m;
    m.q = mc.q;
    m.q2 = mc.q << 1;
    m.qinv64 = mc.qinv64;
    m.qd = mc.qd;
    m.qinvd = mc.qinvd;
    m.lazy = mc.q < kLazyLimit;
    m.recentre = mc.f64_recentre != 0;
    std::vector<u64> sm(N);
    if (!inverse) {
        for (int g = 0; g < BS; g++) {
            typename P::V x[16];
            for (int k = 0; k < 16; k++) x[k] = P::from_raw(data[g + k * BS]);
            fwd_stages<P, 4>(x, tw, 1u, m);
            for (int k = 0; k < 16; k++) sm[k * BS + swz(g)] = P::to_bits(x[k]);
        }
        for (int i = 0; i < PL.nrounds; i++) local_dyn<P, false>(logN, i, PL, sm.data(), tw, m, false, i == PL.nrounds - 1, data);
        (void)sm;
    } else {
        for (int i = PL.nrounds - 1; i >= 0; i--) local_dyn<P, true>(logN, i, PL, sm.data(), tw, m, i == PL.nrounds - 1, false, data);
        for (int g = 0; g < BS; g++) {
            typename P::V x[16];
            for (int k = 0; k < 16; k++) x[k] = P::from_bits(sm[k * BS + swz(g)]);
            inv_stages<P, 4, true>(x, tw, 1u, m, ninv, w_ninv);
            for (int k = 0; k < 16; k++) data[g + k * BS] = P::inv_canonical(x[k], m);
        }
    }
}


// mirrors k_ntt_persist (lsb_ntt_cluster.cu): the same rounds with each block's twiddles taken from its own local
// table (built from the PLAIN table through persist_tw_src, indexed with twbase = 1 and the local tail-SoA stage)
template <class P, int LOGN, bool INV>
static void persist_rounds(u64* sm, const typename P::TW* tws, const ModCtx& m, u64* data) {
    constexpr int N = 1 << LOGN, BS = N / 16, LS = LOGN - 4;
    constexpr Plan PL = make_plan(LS, 0);
    constexpr int LRL = cluster_last_lr(LOGN), S1L = LS - LRL;
    for (int j = 0; j < PL.nrounds; j++) {
        const int i = INV ? PL.nrounds - 1 - j : j;
        const int lr = PL.rounds[i].lr, loggc = plan_loggc(PL, i), v = PL.rounds[i].v;
        const bool last = i == PL.nrounds - 1;
        for (int blk = 0; blk < 16; blk++)
            for (int g = 0; g < (BS >> lr); g++) {
                u64* smt = sm + blk * BS;
                const typename P::TW* tw = tws + (size_t)blk * BS;
                u64* gio = data + blk * BS;
                if (lr == 4 && !last) local_round<P, 4, INV, false, false, -1>(smt, tw, g, loggc, v, 1u, m, gio, gio);
                else if (lr == 3 && !last) local_round<P, 3, INV, false, false, -1>(smt, tw, g, loggc, v, 1u, m, gio, gio);
                else if (lr == 4) local_round<P, 4, INV, INV, !INV, S1L>(smt, tw, g, loggc, v, 1u, m, gio, nullptr);
                else if (lr == 3) local_round<P, 3, INV, INV, !INV, S1L>(smt, tw, g, loggc, v, 1u, m, gio, nullptr);
                else throw std::runtime_error("emul: unhandled persistent round");
            }
    }
}
template <class P, int LOGN>
static void persist_ntt_n(const typename P::TW* plain, const ModConst& mc, typename P::TW ninv, typename P::TW w_ninv, u64* data,
                          bool inverse) {
    constexpr int N = 1 << LOGN, BS = N / 16, LS = LOGN - 4;
    constexpr Plan PL = make_plan(LS, 0);
    constexpr int LRL = cluster_last_lr(LOGN), S1L = LS - LRL;
    ModCtx m;
    m.q = mc.q;
    m.q2 = mc.q << 1;
    m.qinv64 = mc.qinv64;
    m.qd = mc.qd;
    m.qinvd = mc.qinvd;
    m.lazy = mc.q < kLazyLimit;
    m.recentre = mc.f64_recentre != 0;
    std::vector<typename P::TW> tws((size_t)16 * BS), twx(16);
    for (int blk = 0; blk < 16; blk++)
        for (int li = 1; li < BS; li++) tws[(size_t)blk * BS + li] = plain[persist_tw_src((u32)li, blk, LS, LRL)];
    for (int i = 0; i < 16; i++) twx[i] = plain[i];
    std::vector<u64> sm(N);
    if (!inverse) {
        for (int g = 0; g < BS; g++) {
            typename P::V x[16];
            for (int k = 0; k < 16; k++) x[k] = P::from_raw(data[g + k * BS]);
            fwd_stages<P, 4>(x, twx.data(), 1u, m);
            for (int k = 0; k < 16; k++) sm[k * BS + swz(g)] = P::to_bits(x[k]);
        }
        persist_rounds<P, LOGN, false>(sm.data(), tws.data(), m, data);
        for (int c = 0; c < N / 2; c++) {
            const int blk = (2 * c) / BS, u = (2 * c) % BS;
            data[2 * c] = sm[blk * BS + swz(u)];
            data[2 * c + 1] = sm[blk * BS + swz(u) + 1];
        }
    } else {
        persist_rounds<P, LOGN, true>(sm.data(), tws.data(), m, data);
        for (int g = 0; g < BS; g++) {
            typename P::V x[16];
            for (int k = 0; k < 16; k++) x[k] = P::from_bits(sm[k * BS + swz(g)]);
            inv_stages<P, 4, true>(x, twx.data(), 1u, m, ninv, w_ninv);
            for (int k = 0; k < 16; k++) data[g + k * BS] = P::inv_canonical(x[k], m);
        }
    }
}
template <class P>
static void persist_ntt(int logN, const typename P::TW* plain, const ModConst& mc, typename P::TW ninv, typename P::TW w_ninv,
                        u64* data, bool inverse) {
    switch (logN) {
        case 12: return persist_ntt_n<P, 12>(plain, mc, ninv, w_ninv, data, inverse);
        case 13: return persist_ntt_n<P, 13>(plain, mc, ninv, w_ninv, data, inverse);
        case 14: return persist_ntt_n<P, 14>(plain, mc, ninv, w_ninv, data, inverse);
        case 15: return persist_ntt_n<P, 15>(plain, mc, ninv, w_ninv, data, inverse);
        case 16: return persist_ntt_n<P, 16>(plain, mc, ninv, w_ninv, data, inverse);
        default: throw std::runtime_error("emul: unsupported ring degree");
    }
}

extern "C" {

int emul_ntt_persist(int logN, u64 q, u64* data, int inverse, int f64) {
    try {
        const int N = 1 << logN;
        std::vector<Tw> f(N), iv(N);
        std::vector<TwD> fd(N), ivd(N);
        ModConst mc;
        build_ntt_tables(q, logN, f.data(), iv.data(), &mc, fd.data(), ivd.data());
        if (f64) {
            if (!mc.f64_ok) return 2;
            persist_ntt<ArithF64>(logN, inverse ? ivd.data() : fd.data(), mc, mc.ninv_d, mc.w_ninv_d, data, inverse != 0);
        } else {
            persist_ntt<ArithU64>(logN, inverse ? iv.data() : f.data(), mc, mc.ninv, mc.w_ninv, data, inverse != 0);
        }
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "emul_ntt_persist: %s\n", e.what());
        return 1;
    }
}

// f64: 1 = FP64 butterfly policy with (w, w/q) twiddles, 2 = with 8-byte twiddles (both return 2 if the prime does not
// qualify), 0 = 64-bit Shoup/Harvey policy
int emul_ntt_cluster(int logN, u64 q, u64* data, int inverse, int f64) {
    try {
        const int N = 1 << logN;
        std::vector<Tw> f(N), iv(N);
        std::vector<TwD> fd(N), ivd(N);
        ModConst mc;
        build_ntt_tables(q, logN, f.data(), iv.data(), &mc, fd.data(), ivd.data());
        const int lrl = cluster_last_lr(logN);
        soa_tail(f.data(), logN, lrl);
        soa_tail(iv.data(), logN, lrl);
        soa_tail(fd.data(), logN, lrl);
        soa_tail(ivd.data(), logN, lrl);
        if (f64 == 2) {  // FP64 policy with 8-byte twiddles (the device default, LSB_F64_TW8)
            if (!mc.f64_ok) return 2;
            std::vector<double> w(N);
            for (int i = 0; i < N; i++) w[i] = inverse ? ivd[i].w : fd[i].w;
            cluster_ntt<ArithF64W>(logN, w.data(), mc, mc.ninv_d.w, mc.w_ninv_d.w, data, inverse != 0);
        } else if (f64) {
            if (!mc.f64_ok) return 2;
            cluster_ntt<ArithF64>(logN, inverse ? ivd.data() : fd.data(), mc, mc.ninv_d, mc.w_ninv_d, data, inverse != 0);
        } else {
            cluster_ntt<ArithU64>(logN, inverse ? iv.data() : f.data(), mc, mc.ninv, mc.w_ninv, data, inverse != 0);
        }
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "emul_ntt_cluster: %s\n", e.what());
        return 1;
    }
}

int emul_ntt(int logN, u64 q, u64* data, int inverse, int nthreads) {
    try {
        const int N = 1 << logN;
        std::vector<Tw> f(N), iv(N);
        ModConst mc;
        build_ntt_tables(q, logN, f.data(), iv.data(), &mc);
        if (logN <= 14) {
            single_pass(logN, f.data(), iv.data(), mc, data, inverse != 0, nthreads);
        } else if (!inverse) {
            cols_pass(logN, f.data(), mc, data, false);
            rows_pass(logN, f.data(), mc, data, false);
        } else {
            rows_pass(logN, iv.data(), mc, data, true);
            cols_pass(logN, iv.data(), mc, data, true);
        }
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "emul_ntt: %s\n", e.what());
        return 1;
    }
}

// moduli[nmod]; src/dst index lists; in [ns][N], out [nt][N]; mode/centered as BexJob; other [nt][N] (mode 1)
int emul_basis_extend(int logN, const u64* moduli, int nmod, const int* src, int ns, const int* dst, int nt, int centered,
                      const u64* in, u64* out) {
    try {
        const int N = 1 << logN;
        std::vector<u64> mods(moduli, moduli + nmod);
        std::vector<ModConst> mc(nmod);
        for (int m = 0; m < nmod; m++) {
            mc[m].q = mods[m];
            mc[m].bq = host::make_barrett(mods[m]);
            mc[m].qd = (double)mods[m];
            mc[m].qinvd = 1.0 / (double)mods[m];
        }
        BexJob job = make_bex_tables(mods, std::vector<int>(src, src + ns), std::vector<int>(dst, dst + nt));
        job.centered = centered & 1;
        if (centered & 2) job.f64_ok = 0;  // force the integer CRT (primes below 2^47 default to the FP64 one)
        for (int j = 0; j < N; j++) {
            u64 x[kBexMaxSrc];
            for (int i = 0; i < ns; i++) x[i] = in[(size_t)i * N + j];
            bex_coeff(job, mc.data(), x, [&](int) { return (u64)0; }, [&](int t, u64 v) { out[(size_t)t * N + j] = v; });
        }
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "emul_basis_extend: %s\n", e.what());
        return 1;
    }
}

// how many targets of this job the mixed-pipe CRT sends to the FP64 pipe (low byte), whether the sources are split (bit 8),
// whether the whole job runs on the FP64 pipe (bit 9)
int emul_bex_mixed_targets(const u64* moduli, int nmod, const int* src, int ns, const int* dst, int nt) {
    std::vector<u64> mods(moduli, moduli + nmod);
    BexJob job = make_bex_tables(mods, std::vector<int>(src, src + ns), std::vector<int>(dst, dst + nt));
    return job.n_f64 | (job.src_split << 8) | (job.f64_ok << 9);
}

// a*b mod q through the device Barrett path, and 128-bit accumulate path
u64 emul_mulmod(u64 a, u64 b, u64 q) { return mulmod(a, b, host::make_barrett(q)); }
u64 emul_reduce_wide(u64 lo, u64 hi, u64 q) { return reduce_wide(lo, hi, host::make_barrett(q)); }
u64 emul_barrett64(u64 a, u64 q) { return barrett64(a, host::make_barrett(q)); }
u64 emul_mul_shoup(u64 a, u64 w, u64 q) { return mul_shoup(a, host::make_tw(w, q), q); }

}  // extern "C"
