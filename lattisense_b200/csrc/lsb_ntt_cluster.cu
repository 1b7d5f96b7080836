// Cluster NTT / INTT for sm_100a: one thread-block CLUSTER per limb, the limb split over the CTAs' shared
// memory (32-64 KiB each), the cross-CTA radix-16 round exchanged through distributed shared memory.
//
// Why: a whole N=2^14 limb is 128 KiB, so the single-CTA kernel (lsb_ntt.cu) runs one CTA per SM and its
// load / butterfly / store phases never overlap (ncu: 25 % warps active, profiles/r01_ntt14_ncu_full_summary.txt).
// Here the first forward round (stages 0-3, gaps N/2..N/16) reads global memory directly and scatters its
// outputs to the CTA that owns the 1/16-block they fall in; after one cluster barrier every CTA finishes its
// blocks (independent N/16-point sub-transforms) locally.  Several clusters are resident per SM in different
// phases, one HBM read + one HBM write per limb is kept (16*N bytes), and N = 2^15, 2^16 no longer need the
// two-pass kernels (32*N bytes).  The inverse runs the same rounds backwards; its last round gathers from the
// peers' shared memory.  Butterfly arithmetic is a policy (lsb_ntt_arith.h): FP64 for primes below ~2^47,
// 64-bit Shoup/Harvey otherwise; both give canonical residues identical to the oracle (oracle/orc_core.hpp:55-94,
// Lattigo ring.NTT / ring.InvNTT ordering).
#include <cooperative_groups.h>
#include <cstdlib>
#include <map>
#include <mutex>
#include <type_traits>

#include "lsb_kernels.h"
#include "lsb_ntt_core.h"
#include "lsb_ntt_cluster_core.h"
#include "lsb_ntt_persist_core.h"

// N = 2^16: 8 CTAs per limb hold 64 KiB tiles (3 CTAs = 24 warps per SM; ncu: 31 % warps active, long-scoreboard bound,
// profiles/r02_ks16_integer_ncu.txt); 16 CTAs per limb (non-portable cluster size) halve the tile: measured on the
// level-24 key-switch (60/61-bit primes) 6.11 -> 5.57 ms at batch 16, 0.606 -> 0.491 ms for a single ciphertext (4 CTAs
// per SM asked of ptxas; 3 / 5 / 6 were slower)
#ifndef LSB_CL16
#define LSB_CL16 16
#endif
#ifndef LSB_CL16_MINB
#define LSB_CL16_MINB 4
#endif
#ifndef LSB_FIN_PREFETCH
#define LSB_FIN_PREFETCH 0  // measured: +0.6 GB of DRAM reads per step (prefetched lines evicted before use), no time gain
#endif
#ifndef LSB_CL_RMAJOR
#define LSB_CL_RMAJOR 0
#endif
#ifndef LSB_CL_MINB
#define LSB_CL_MINB 4
#endif
#ifndef LSB_CL_NT
#define LSB_CL_NT 256
#endif
// N = 2^14 (the headline ring): 8 CTAs of 128 threads and 16 KiB per limb, 8 resident per SM.  Measured against the
// other shapes with the same 32 warps per SM (tools/ab_ntt.sh): 4 CTAs x 256 threads 2599 / 2321 GB/s (fwd / inv),
// 2 x 256 at 3 per SM 2767 / 2466, 8 x 128 2745 / 2492.
#ifndef LSB_CL14
#define LSB_CL14 8
#endif
#ifndef LSB_CL14_NT
#define LSB_CL14_NT 128
#endif
#ifndef LSB_CL14_MINB
#define LSB_CL14_MINB 8
#endif
// forward transforms at N = 2^14 ask ptxas for 9 resident CTAs (56 registers): with the 8-byte twiddles the forward kernel
// fits them without spills and gains 4 % (2.92 -> 3.03 TB/s), the inverse does not (2.71 -> 2.70), 10 loses 4 %
// resident CTAs asked of ptxas for the FP64 policy at N = 2^16 / 2^15 (tools/ab_f64_16.sh; measured with the 8-byte twiddles:
// 4 / 5 / 6 -> forward 2219 / 2276 / 2169 GB/s, inverse 1870 / 1820 / 1606 GB/s at N = 2^16, 2652 / 2589 / 2113 and
// 2314 / 2142 / 1223 at N = 2^15: the defaults stay)
#ifndef LSB_CL16_MINB_F64
#define LSB_CL16_MINB_F64 LSB_CL16_MINB
#endif
#ifndef LSB_CL15_MINB_F64
#define LSB_CL15_MINB_F64 LSB_CL_MINB
#endif
#ifndef LSB_CL14_MINB_FWD
#define LSB_CL14_MINB_FWD 9
#endif
// FP64 policy of the cluster kernel: 1 = 8-byte twiddles (ArithF64W), 0 = (w, w/q) pairs (ArithF64)
#ifndef LSB_F64_TW8
#define LSB_F64_TW8 1
#endif
namespace cg = cooperative_groups;

namespace lsb {

template <int LOGN>
struct ClusterCfg {
    static constexpr int N = 1 << LOGN;
    static constexpr int CL = LOGN <= 12 ? 1 : LOGN == 13 ? 2 : LOGN == 14 ? LSB_CL14 : LOGN == 16 ? LSB_CL16 : 8;  // CTAs per limb
    static constexpr int BS = N / 16;                                                 // block = sub-transform size
    static constexpr int LS = LOGN - 4;
    static constexpr int NB = 16 / CL;      // blocks per CTA
    static constexpr int TILE = N / CL;     // words of shared memory per CTA
    static constexpr int NT = LOGN == 14 ? LSB_CL14_NT : LSB_CL_NT;
    // resident CTAs per SM asked of ptxas; N = 2^16 tiles are 64 KiB, so only 3 fit anyway and asking for 3 leaves the
    // compiler 85 registers (inverse: 1365 -> 1538 GB/s on the mixed chain, 1004 -> 1098 GB/s on 61-bit primes)
    static constexpr int MINB = LOGN == 14 ? LSB_CL14_MINB : LOGN == 16 ? LSB_CL16_MINB : LSB_CL_MINB;
    static constexpr int G0 = BS / CL;      // cross-round groups per CTA
    // f64w: the FP64 policy with 8-byte twiddles (the integer policy spills at 56 registers)
    static constexpr int minb(bool inv, bool f64w) {
        return LOGN == 14 && !inv && f64w ? LSB_CL14_MINB_FWD : LOGN == 16 && f64w ? LSB_CL16_MINB_F64 : LOGN == 15 && f64w ? LSB_CL15_MINB_F64 : MINB;
    }
};

constexpr int kMaxPeers = 8;
#if LSB_F64_TW8
typedef ArithF64W ArithF64C;
#else
typedef ArithF64 ArithF64C;
#endif

struct NttClArgs {
    BatchPtr in, out;
    int total_rows;
    const Tw* tw;     // [nmod][N]
    const TwD* tw_d;  // [nmod][N]
    const double* tw_w;  // [nmod][N], w only (ArithF64W)
    const ModConst* mods;
    RowMap map;
    // fused finish epilogue (forward only): out[fm.out_slot[f]] = (fa[fm.a_slot[f]] - NTT(row)) * fm.c[f] (+ fadd[fm.add_slot[f]])
    // with f = map.out_slot[row]; the transform itself is not stored
    const FinishMap* fm;
    BatchPtr fa, fadd, fout;
    int has_add;
    // fused all-gather (inverse only, limb-sharded key-switch): instead of `out`, row r is stored at word offset
    // out_slot[r] * N of EVERY buffer in peer[] (this rank's own copy included) -- peer-mapped device memory of the
    // other ranks' GPUs, written over NVLink while the transform's remaining groups are still being computed
    int npeer;
    u64* peer[kMaxPeers];
    // persistent kernel (k_ntt_persist): limbs are visited row-major (work item i -> row i / batch, item i % batch) so
    // that a cluster stays on one modulus; twp / twp_d are the PLAIN bit-reversed tables it stages into shared memory
    int batch;
    const Tw* twp;
    const TwD* twp_d;
};

template <int I, int NR, class F>
__device__ __forceinline__ void static_for_cl(F&& f) {
    if constexpr (I < NR) {
        f(std::integral_constant<int, I>{});
        static_for_cl<I + 1, NR>(f);
    }
}

template <class P>
__device__ __forceinline__ const typename P::TW* tw_of(const NttClArgs& a, int mod, int N) {
    if constexpr (std::is_same<P, ArithF64W>::value) return a.tw_w + (size_t)mod * N;
    else if constexpr (std::is_same<P, ArithF64>::value) return a.tw_d + (size_t)mod * N;
    else return a.tw + (size_t)mod * N;
}

// Inverse transform: the trailing cluster barrier ("nobody reads my tile any more") is split -- a thread arrives as soon as the
// first stage of its cross round has consumed every word it loaded from the peers' tiles (so the loads have completed), and
// waits only after the remaining stages and the global stores.  0: one cluster.sync() at the end.
#ifndef LSB_INV_SPLIT_BARRIER
#define LSB_INV_SPLIT_BARRIER 1
#endif
__device__ __forceinline__ void cl_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cl_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

template <int LOGN, bool INV, class P, bool FIN, bool PEERS = false>
__device__ __forceinline__ void ntt_cluster_body(const NttClArgs& a, u64* sm) {
    typedef ClusterCfg<LOGN> C;
    constexpr int N = C::N, CL = C::CL, BS = C::BS, NB = C::NB, TILE = C::TILE, NT = C::NT;
    constexpr Plan PL = make_plan(C::LS, 0);
    const int tid = threadIdx.x;
    const int rank = CL > 1 ? (int)(blockIdx.x % CL) : 0;
    const int row = blockIdx.x / CL;
#if LSB_CL_RMAJOR
    // modulus-major order: consecutive clusters transform the same limb of consecutive batch items, so the CTAs that
    // share an SM read the same twiddles (L1 hits instead of L2 round trips)
    const int r = row / a.batch, b = row - r * a.batch;
#else
    const int b = row / a.map.n, r = row - b * a.map.n;
#endif
    const u64* src = a.in.at(b) + (size_t)a.map.in_slot[r] * N;
    u64* dst = a.out.at(b) + (size_t)a.map.out_slot[r] * N;
    const int mod = a.map.mod[r];
    const ModConst& mc = a.mods[mod];
    const typename P::TW* tw = tw_of<P>(a, mod, N);
    ModCtx m;
    m.q = mc.q;
    m.q2 = mc.q << 1;
    m.qinv64 = mc.qinv64;
    m.qd = mc.qd;
    m.qinvd = mc.qinvd;
    m.lazy = mc.q < kLazyLimit;
    m.recentre = mc.f64_recentre != 0;
    cg::cluster_group cluster = cg::this_cluster();

    if (!INV) {
#if LSB_FIN_PREFETCH
        if constexpr (FIN) {
            // the finish epilogue reads this CTA's slice of `a` (and of the addend) only after every round: start those
            // HBM reads now, into L2, so that the epilogue's loads find them there (no registers, no shared memory)
            const int f = a.map.out_slot[r];
            const FinishMap& fm = *a.fm;
            const char* pa = reinterpret_cast<const char*>(a.fa.at(b) + (size_t)fm.a_slot[f] * N + (size_t)rank * TILE);
            for (int c = tid * 128; c < TILE * 8; c += NT * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(pa + c));
            if (a.has_add && fm.add_slot[f] != 0xFF) {
                const char* pz = reinterpret_cast<const char*>(a.fadd.at(b) + (size_t)fm.add_slot[f] * N + (size_t)rank * TILE);
                for (int c = tid * 128; c < TILE * 8; c += NT * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(pz + c));
            }
        }
#endif
        // ---- cross round: stages 0..3 on words g + k*BS, outputs scattered to the owners of block k ----
        u64* peer[CL];
#pragma unroll
        for (int c = 0; c < CL; c++) peer[c] = CL > 1 ? cluster.map_shared_rank(sm, c) : sm;
        for (int gl = tid; gl < C::G0; gl += NT) {
            const int g = rank * C::G0 + gl;
            typename P::V x[16];
#pragma unroll
            for (int k = 0; k < 16; k++) x[k] = P::from_raw(src[g + k * BS]);
            fwd_stages<P, 4>(x, tw, 1u, m);
#pragma unroll
            for (int k = 0; k < 16; k++) peer[k / NB][(k % NB) * BS + swz(g)] = P::to_bits(x[k]);
        }
        if (CL > 1) cluster.sync();
        else __syncthreads();
        // ---- local rounds on the NB blocks this CTA owns ----
        static_for_cl<0, PL.nrounds>([&](auto ic) {
            constexpr int i = decltype(ic)::value;
            constexpr int lr = PL.rounds[i].lr;
            constexpr int loggc = plan_loggc(PL, i);
            constexpr bool last = i == PL.nrounds - 1;
            constexpr int GPB = BS >> lr;  // groups per block
            for (int gg = tid; gg < NB * GPB; gg += NT) {
                const int st = gg / GPB, g = gg - st * GPB;
                local_round<P, lr, false, false, last, (last ? LOGN - lr : -1)>(sm + st * BS, tw, g, loggc, PL.rounds[i].v,
                                                                                16u + (u32)(rank * NB + st), m, nullptr, nullptr);
            }
            __syncthreads();
        });
        if constexpr (!FIN) {
            u64* out = dst + (size_t)rank * TILE;
            for (int c = tid; c < TILE / 2; c += NT) reinterpret_cast<W2*>(out)[c] = *reinterpret_cast<const W2*>(&sm[swz(2 * c)]);
        } else {
            // ModDown / rescale finish on the integer pipe while other CTAs of the SM run FP64 butterflies
            const int f = a.map.out_slot[r];
            const FinishMap& fm = *a.fm;
            const Tw cmul = fm.c[f];
            const u64 q = mc.q;
            const size_t off = (size_t)rank * (TILE / 2);
            const W2* pa = reinterpret_cast<const W2*>(a.fa.at(b) + (size_t)fm.a_slot[f] * N) + off;
            const bool add = a.has_add && fm.add_slot[f] != 0xFF;
            const W2* pz = add ? reinterpret_cast<const W2*>(a.fadd.at(b) + (size_t)fm.add_slot[f] * N) + off : nullptr;
            W2* po = reinterpret_cast<W2*>(a.fout.at(b) + (size_t)fm.out_slot[f] * N) + off;
            for (int c = tid; c < TILE / 2; c += NT) {
                const W2 y = *reinterpret_cast<const W2*>(&sm[swz(2 * c)]);
                const W2 x = pa[c];
                W2 o;
                o.a = mul_shoup(submod(x.a, y.a, q), cmul, q);
                o.b = mul_shoup(submod(x.b, y.b, q), cmul, q);
                if (add) {
                    const W2 z = pz[c];
                    o.a = addmod(o.a, z.a, q);
                    o.b = addmod(o.b, z.b, q);
                }
                po[c] = o;
            }
        }
    } else {
        static_for_cl<0, PL.nrounds>([&](auto ic) {
            constexpr int i = PL.nrounds - 1 - decltype(ic)::value;
            constexpr int lr = PL.rounds[i].lr;
            constexpr int loggc = plan_loggc(PL, i);
            constexpr bool first = i == PL.nrounds - 1;
            constexpr int GPB = BS >> lr;
            for (int gg = tid; gg < NB * GPB; gg += NT) {
                const int st = gg / GPB, g = gg - st * GPB;
                local_round<P, lr, true, first, false, (first ? LOGN - lr : -1)>(sm + st * BS, tw, g, loggc, PL.rounds[i].v,
                                                                                 16u + (u32)(rank * NB + st), m,
                                                                                 src + (size_t)rank * TILE + st * BS);
            }
            __syncthreads();
        });
        if (CL > 1) cluster.sync();
        const u64* peer[CL];
#pragma unroll
        for (int c = 0; c < CL; c++) peer[c] = CL > 1 ? cluster.map_shared_rank(sm, c) : sm;
        typename P::TW ninv, w_ninv;
        if constexpr (std::is_same<P, ArithF64W>::value) {
            ninv = mc.ninv_d.w;
            w_ninv = mc.w_ninv_d.w;
        } else if constexpr (std::is_same<P, ArithF64>::value) {
            ninv = mc.ninv_d;
            w_ninv = mc.w_ninv_d;
        } else {
            ninv = mc.ninv;
            w_ninv = mc.w_ninv;
        }
        // (every thread runs the loop body exactly once when G0 == NT, so the barrier instructions inside it are convergent)
        constexpr bool SPLIT = LSB_INV_SPLIT_BARRIER && CL > 1 && C::G0 == NT;
        for (int gl = tid; gl < C::G0; gl += NT) {
            const int g = rank * C::G0 + gl;
            typename P::V x[16];
#pragma unroll
            for (int k = 0; k < 16; k++) x[k] = P::from_bits(peer[k / NB][(k % NB) * BS + swz(g)]);
            if constexpr (SPLIT) {
                inv_stages<P, 4, true, -1, 3, 3>(x, tw, 1u, m, ninv, w_ninv);  // 8 butterflies over all 16 loaded words
                cl_arrive_relaxed();
                inv_stages<P, 4, true, -1, 2, 0>(x, tw, 1u, m, ninv, w_ninv);
            } else {
                inv_stages<P, 4, true>(x, tw, 1u, m, ninv, w_ninv);
            }
            if constexpr (!PEERS) {
#pragma unroll
                for (int k = 0; k < 16; k++) dst[g + k * BS] = P::inv_canonical(x[k], m);
            } else {
                const size_t off = (size_t)a.map.out_slot[r] * N + g;
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    const u64 v = P::inv_canonical(x[k], m);
                    for (int pr = 0; pr < a.npeer; pr++) a.peer[pr][off + k * BS] = v;
                }
            }
        }
        if constexpr (SPLIT) cl_wait();  // every thread of the cluster has arrived: nobody reads this CTA's tile any more
        else if (CL > 1) cluster.sync();  // peers may still be reading this CTA's shared memory
    }
}

template <int LOGN, bool INV, class P, bool FIN = false, bool PEERS = false>
__global__ void __launch_bounds__(ClusterCfg<LOGN>::NT, ClusterCfg<LOGN>::minb(INV, std::is_same<P, ArithF64W>::value)) k_ntt_cluster(const __grid_constant__ NttClArgs a) {
    extern __shared__ __align__(1024) u64 sm[];
    ntt_cluster_body<LOGN, INV, P, FIN, PEERS>(a, sm);
}

// ---------------------------------------------------------------------------------------------------------------
// Persistent variant (N <= 2^14; OPT-IN with LSB_NTT_PERSIST=1, measured SLOWER than one cluster per limb on B200:
// 2.33 vs 2.75 TB/s forward, 2.15 vs 2.49 inverse at N = 2^14 -- profiles/r02_ntt_persist_experiment.txt; it removes
// every long-scoreboard stall but its 48 KiB of shared memory per CTA leave 16 warps per SM instead of 32, and the
// FP64 dependency latency then shows as stall_wait / short_scoreboard): a fixed set of clusters, as many as the device holds at once, each walking
// a contiguous range of the launch's limbs in row-major order (all limbs of one modulus before the next).  What this
// buys over one cluster per limb (ncu of k_ntt_cluster, profiles/r02_ntt14_base_stalls.txt: a third of all warp
// stalls are long-scoreboard waits, ~45 % of them on twiddle loads from L2 and ~35 % on the transform's own input):
//   * the twiddles of the blocks a CTA owns (2^LS - 1 per block; as many bytes as the data tile itself) are staged
//     in SHARED memory once per modulus and re-used for every limb of that modulus -- the local rounds issue no
//     global load at all;
//   * the next limb's input is loaded into registers right after the current limb's cross round has been handed to
//     the peers, i.e. a whole limb ahead of its use: its HBM latency hides behind the cluster barrier and the local
//     rounds.  Occupancy is set by shared memory (48 KiB per CTA at N = 2^14: 4 CTAs per SM), so the 32 extra
//     registers are free;
//   * the cluster barrier is split: "my tile may be overwritten" is signalled (arrive) after the output copy and
//     awaited only after the next limb's cross-round butterflies, "scatter done" is signalled before the prefetch is
//     issued and awaited after it.
// Arithmetic, indexing and data layout are those of k_ntt_cluster (same round code), so the limbs are identical.
template <int LOGN>
struct PersistCfg {
    typedef ClusterCfg<LOGN> C;
    static constexpr int NT = LOGN == 14 ? 128 : 256;
    static constexpr int MINB = LOGN == 14 ? 4 : 2;
    static constexpr int LRL = cluster_last_lr(LOGN);
    static constexpr int S1L = C::LS - LRL;           // local first stage of the contiguous last round
    static constexpr int GPT = C::G0 / NT;            // cross-round groups per thread
    static constexpr size_t SMEM = (size_t)C::TILE * 8 + (size_t)C::NB * C::BS * 16 + 16 * 16 + 16;  // + mbarrier
    static_assert(C::G0 % NT == 0, "cross-round groups must divide evenly over the CTA");
};

__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// "I no longer READ shared memory of this phase": no data is published, so no release fence (which would also wait for
// the output copy's global stores to drain -- 10 % of all stall samples in profiles/r02_ntt14_persist_v1_stalls.txt)
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }

// ---- asynchronous scatter (forward cross round): st.async writes one word into a peer's tile and counts its bytes on
// the PEER's mbarrier, so the receiver just waits for TILE * 8 bytes to land -- no release fence after the remote
// stores (ERRBAR: 9 % of stall samples) and no cluster-wide barrier for this phase
#ifndef LSB_PERSIST_STASYNC
#define LSB_PERSIST_STASYNC 1
#endif
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ u32 mapa_u32(u32 local, int rank) {
    u32 r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_async_u64(u32 remote_addr, u64 v, u32 remote_mbar) {
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr), "l"(v), "r"(remote_mbar)
                 : "memory");
}
__device__ __forceinline__ void mbar_init(u32 mbar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u32 mbar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(u32 mbar, u32 parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "LSB_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra LSB_MBAR_DONE;\n"
        "bra LSB_MBAR_WAIT;\n"
        "LSB_MBAR_DONE:\n"
        "}\n" ::"r"(mbar),
        "r"(parity)
        : "memory");
}

template <class P>
__device__ __forceinline__ const typename P::TW* twp_of(const NttClArgs& a, int mod, int N) {
    if constexpr (std::is_same<P, ArithF64>::value) return a.twp_d + (size_t)mod * N;
    else return a.twp + (size_t)mod * N;
}

template <int LOGN, bool INV, class P, bool FIN>
__global__ void __launch_bounds__(PersistCfg<LOGN>::NT, PersistCfg<LOGN>::MINB) k_ntt_persist(const __grid_constant__ NttClArgs a) {
    typedef ClusterCfg<LOGN> C;
    typedef PersistCfg<LOGN> PC;
    typedef typename P::TW TW;
    constexpr int N = C::N, CL = C::CL, BS = C::BS, NB = C::NB, TILE = C::TILE, NT = PC::NT, GPT = PC::GPT;
    constexpr Plan PL = make_plan(C::LS, 0);
    extern __shared__ __align__(1024) u64 sm[];
    TW* tws = reinterpret_cast<TW*>(sm + TILE);  // [NB][BS] local tables (lsb_ntt_persist_core.h)
    TW* twx = tws + NB * BS;                     // [16] cross-round twiddles (entries 1..15 of the plain table)
    const int tid = threadIdx.x;
    const int rank = CL > 1 ? (int)(blockIdx.x % CL) : 0;
    const int ncl = (int)(gridDim.x / CL), cid = (int)(blockIdx.x / CL);
    const int per = (a.total_rows + ncl - 1) / ncl;
    const int i0 = cid * per, i1 = min(a.total_rows, i0 + per);
    cg::cluster_group cluster = cg::this_cluster();
    u64* peer[CL];
#pragma unroll
    for (int c = 0; c < CL; c++) peer[c] = CL > 1 ? cluster.map_shared_rank(sm, c) : sm;
    constexpr bool ASYNC = LSB_PERSIST_STASYNC && CL > 1 && !INV;
    const u32 mbar = smem_u32(twx + 16);
    u32 phase = 0;
    if (ASYNC && tid == 0) mbar_init(mbar, 1);
    if (CL > 1) cluster.sync();  // every CTA of the cluster is resident (and its mbarrier initialised) before it is addressed
    if (i0 >= i1) return;        // (uniform over the cluster)

    int cur_mod = -1;
    ModCtx m{};
    auto row_of = [&](int i, int& b, int& r) {
        r = i / a.batch;
        b = i - r * a.batch;
    };
    auto stage_twiddles = [&](int mod) {
        const TW* plain = twp_of<P>(a, mod, N);
        for (int idx = tid; idx < NB * BS; idx += NT) {
            const int st = idx / BS, li = idx - st * BS;
            if (li) tws[idx] = plain[persist_tw_src((u32)li, rank * NB + st, C::LS, PC::LRL)];
        }
        if (tid < 16) twx[tid] = plain[tid];
        const ModConst& mc = a.mods[mod];
        m.q = mc.q;
        m.q2 = mc.q << 1;
        m.qinv64 = mc.qinv64;
        m.qd = mc.qd;
        m.qinvd = mc.qinvd;
        m.lazy = mc.q < kLazyLimit;
        m.recentre = mc.f64_recentre != 0;
        __syncthreads();
    };

    if (!INV) {
        u64 xn[GPT][16];  // next limb's cross-round input, raw residues
        auto prefetch = [&](int i) {
            int b, r;
            row_of(i, b, r);
            const u64* src = a.in.at(b) + (size_t)a.map.in_slot[r] * N;
#pragma unroll
            for (int j = 0; j < GPT; j++) {
                const int g = rank * C::G0 + tid + j * NT;
#pragma unroll
                for (int k = 0; k < 16; k++) xn[j][k] = src[g + k * BS];
            }
        };
        prefetch(i0);
        for (int i = i0; i < i1; i++) {
            int b, r;
            row_of(i, b, r);
            const int mod = a.map.mod[r];
            if (mod != cur_mod) {
                stage_twiddles(mod);
                cur_mod = mod;
            }
            // ---- cross round on the prefetched words g + k*BS ----
            typename P::V x[GPT][16];
#pragma unroll
            for (int j = 0; j < GPT; j++) {
#pragma unroll
                for (int k = 0; k < 16; k++) x[j][k] = P::from_raw(xn[j][k]);
                fwd_stages<P, 4>(x[j], twx, 1u, m);
            }
            if (CL > 1 && i > i0) cluster_wait();  // every CTA of the cluster is done with the previous limb's tile
            else if (CL == 1) __syncthreads();
            if constexpr (ASYNC) {
                if (tid == 0) mbar_expect_tx(mbar, TILE * 8);
                const u32 tile0 = smem_u32(sm);
#pragma unroll
                for (int j = 0; j < GPT; j++) {
                    const int g = rank * C::G0 + tid + j * NT;
                    const u32 off = (u32)swz(g) * 8;
#pragma unroll
                    for (int c = 0; c < CL; c++) {
                        const u32 rt = mapa_u32(tile0, c), rm = mapa_u32(mbar, c);
#pragma unroll
                        for (int kk = 0; kk < NB; kk++) st_async_u64(rt + (u32)(kk * BS) * 8 + off, P::to_bits(x[j][c * NB + kk]), rm);
                    }
                }
                if (i + 1 < i1) prefetch(i + 1);
                mbar_wait(mbar, phase);
                phase ^= 1;
            } else {
#pragma unroll
                for (int j = 0; j < GPT; j++) {
                    const int g = rank * C::G0 + tid + j * NT;
#pragma unroll
                    for (int k = 0; k < 16; k++) peer[k / NB][(k % NB) * BS + swz(g)] = P::to_bits(x[j][k]);
                }
                if (CL > 1) cluster_arrive();
                if (i + 1 < i1) prefetch(i + 1);
                if (CL > 1) cluster_wait();
                else __syncthreads();
            }
            // ---- local rounds on the NB blocks this CTA owns, twiddles from shared memory ----
            static_for_cl<0, PL.nrounds>([&](auto ic) {
                constexpr int ir = decltype(ic)::value;
                constexpr int lr = PL.rounds[ir].lr;
                constexpr int loggc = plan_loggc(PL, ir);
                constexpr bool last = ir == PL.nrounds - 1;
                constexpr int GPB = BS >> lr;
#pragma unroll
                for (int gg = tid; gg < NB * GPB; gg += NT) {
                    const int st = gg / GPB, g = gg - st * GPB;
                    local_round<P, lr, false, false, last, (last ? PC::S1L : -1)>(sm + st * BS, tws + st * BS, g, loggc,
                                                                                   PL.rounds[ir].v, 1u, m, nullptr, nullptr);
                }
                __syncthreads();
            });
            if constexpr (!FIN) {
                u64* out = a.out.at(b) + (size_t)a.map.out_slot[r] * N + (size_t)rank * TILE;
                for (int c = tid; c < TILE / 2; c += NT) reinterpret_cast<W2*>(out)[c] = *reinterpret_cast<const W2*>(&sm[swz(2 * c)]);
            } else {
                const int f = a.map.out_slot[r];
                const FinishMap& fm = *a.fm;
                const Tw cmul = fm.c[f];
                const u64 q = m.q;
                const size_t off = (size_t)rank * (TILE / 2);
                const W2* pa = reinterpret_cast<const W2*>(a.fa.at(b) + (size_t)fm.a_slot[f] * N) + off;
                const bool add = a.has_add && fm.add_slot[f] != 0xFF;
                const W2* pz = add ? reinterpret_cast<const W2*>(a.fadd.at(b) + (size_t)fm.add_slot[f] * N) + off : nullptr;
                W2* po = reinterpret_cast<W2*>(a.fout.at(b) + (size_t)fm.out_slot[f] * N) + off;
                for (int c = tid; c < TILE / 2; c += NT) {
                    const W2 y = *reinterpret_cast<const W2*>(&sm[swz(2 * c)]);
                    const W2 xa = pa[c];
                    W2 o;
                    o.a = mul_shoup(submod(xa.a, y.a, q), cmul, q);
                    o.b = mul_shoup(submod(xa.b, y.b, q), cmul, q);
                    if (add) {
                        const W2 z = pz[c];
                        o.a = addmod(o.a, z.a, q);
                        o.b = addmod(o.b, z.b, q);
                    }
                    po[c] = o;
                }
            }
            if (CL > 1 && i + 1 < i1) cluster_arrive_relaxed();  // this thread no longer reads the tile
        }
    } else {
        constexpr int LRL = PC::LRL, GPBL = BS >> LRL, RL = 1 << LRL;
        constexpr int GPTL = (NB * GPBL) / NT;  // first-inverse-round groups per thread
        static_assert((NB * GPBL) % NT == 0, "first inverse round must divide evenly over the CTA");
        W2 xn[GPTL][RL / 2];
        auto prefetch = [&](int i) {
            int b, r;
            row_of(i, b, r);
            const u64* src = a.in.at(b) + (size_t)a.map.in_slot[r] * N + (size_t)rank * TILE;
#pragma unroll
            for (int j = 0; j < GPTL; j++) {
                const int gg = tid + j * NT;
                const W2* p = reinterpret_cast<const W2*>(src + (size_t)gg * RL);  // st * BS + g * RL == gg * RL
#pragma unroll
                for (int k = 0; k < RL / 2; k++) xn[j][k] = p[k];
            }
        };
        prefetch(i0);
        for (int i = i0; i < i1; i++) {
            int b, r;
            row_of(i, b, r);
            const int mod = a.map.mod[r];
            if (mod != cur_mod) {
                stage_twiddles(mod);
                cur_mod = mod;
            }
            u64* dst = a.out.at(b) + (size_t)a.map.out_slot[r] * N;
            // ---- first inverse round (contiguous) from the prefetched registers into the tile ----
            if (CL > 1 && i > i0) cluster_wait();  // the peers have gathered the previous limb out of this tile
#pragma unroll
            for (int j = 0; j < GPTL; j++) {
                const int gg = tid + j * NT;
                const int st = gg / GPBL, g = gg - st * GPBL;
                typename P::V x[RL];
#pragma unroll
                for (int k = 0; k < RL / 2; k++) {
                    x[2 * k] = P::from_raw(xn[j][k].a);
                    x[2 * k + 1] = P::from_raw(xn[j][k].b);
                }
                inv_stages<P, LRL, false, PC::S1L>(x, tws + st * BS, ((u32)1 << PC::S1L) + (u32)g, m, TW{}, TW{});
                u64* smt = sm + st * BS;
#pragma unroll
                for (int k = 0; k < RL; k += 2) {
                    W2 pr;
                    pr.a = P::to_bits(x[k]);
                    pr.b = P::to_bits(x[k + 1]);
                    *reinterpret_cast<W2*>(&smt[swz(g * RL + k)]) = pr;
                }
            }
            __syncthreads();
            if (i + 1 < i1) prefetch(i + 1);
            static_for_cl<1, PL.nrounds>([&](auto ic) {
                constexpr int ir = PL.nrounds - 1 - decltype(ic)::value;
                constexpr int lr = PL.rounds[ir].lr;
                constexpr int loggc = plan_loggc(PL, ir);
                constexpr int GPB = BS >> lr;
#pragma unroll
                for (int gg = tid; gg < NB * GPB; gg += NT) {
                    const int st = gg / GPB, g = gg - st * GPB;
                    local_round<P, lr, true, false, false, -1>(sm + st * BS, tws + st * BS, g, loggc, PL.rounds[ir].v, 1u, m,
                                                               nullptr, nullptr);
                }
                __syncthreads();
            });
            if (CL > 1) {
                cluster_arrive();
                cluster_wait();
            }
            TW ninv, w_ninv;
            {
                const ModConst& mc = a.mods[mod];
                if constexpr (std::is_same<P, ArithF64>::value) {
                    ninv = mc.ninv_d;
                    w_ninv = mc.w_ninv_d;
                } else {
                    ninv = mc.ninv;
                    w_ninv = mc.w_ninv;
                }
            }
#pragma unroll
            for (int j = 0; j < GPT; j++) {
                const int g = rank * C::G0 + tid + j * NT;
                typename P::V x[16];
#pragma unroll
                for (int k = 0; k < 16; k++) x[k] = P::from_bits(peer[k / NB][(k % NB) * BS + swz(g)]);
                inv_stages<P, 4, true>(x, twx, 1u, m, ninv, w_ninv);
#pragma unroll
                for (int k = 0; k < 16; k++) dst[g + k * BS] = P::inv_canonical(x[k], m);
            }
            if (CL > 1 && i + 1 < i1) cluster_arrive_relaxed();  // done reading the peers' tiles
            else if (CL == 1) __syncthreads();
        }
        if (CL > 1) cluster.sync();  // no CTA leaves while a peer may still read its shared memory
    }
}

template <int LOGN, class K>
static void launch_persist_k(K kernel, NttClArgs a, int sm_count, cudaStream_t s) {
    typedef ClusterCfg<LOGN> C;
    typedef PersistCfg<LOGN> PC;
    // per kernel (K is the same function-pointer type for every instantiation, so key by address)
    static std::mutex mu;
    static std::map<const void*, int> cache;
    std::lock_guard<std::mutex> lock(mu);
    int& max_clusters = cache[(const void*)kernel];
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(PC::NT, 1, 1);
    cfg.dynamicSmemBytes = PC::SMEM;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C::CL;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (max_clusters == 0) {
        LSB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PC::SMEM));
        LSB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        cfg.gridDim = dim3((unsigned)(sm_count * PC::MINB), 1, 1);
        int n = 0;
        LSB_CUDA(cudaOccupancyMaxActiveClusters(&n, kernel, &cfg));
        static const int cap = [] {
            const char* e = getenv("LSB_NTT_PERSIST_CLUSTERS");
            return e ? atoi(e) : 0;
        }();
        max_clusters = std::max(1, cap > 0 ? std::min(cap, n) : n);
    }
    const int ncl = std::min(a.total_rows, max_clusters);
    cfg.gridDim = dim3((unsigned)(ncl * C::CL), 1, 1);
    LSB_CUDA(cudaLaunchKernelEx(&cfg, kernel, a));
}

template <int LOGN>
static void launch_persist_n(const NttClArgs& a, bool inverse, bool f64, int sm_count, cudaStream_t s) {
    if (a.fm) {
        if (f64) launch_persist_k<LOGN>(k_ntt_persist<LOGN, false, ArithF64, true>, a, sm_count, s);
        else launch_persist_k<LOGN>(k_ntt_persist<LOGN, false, ArithU64, true>, a, sm_count, s);
        return;
    }
    if (f64) {
        if (inverse) launch_persist_k<LOGN>(k_ntt_persist<LOGN, true, ArithF64, false>, a, sm_count, s);
        else launch_persist_k<LOGN>(k_ntt_persist<LOGN, false, ArithF64, false>, a, sm_count, s);
    } else {
        if (inverse) launch_persist_k<LOGN>(k_ntt_persist<LOGN, true, ArithU64, false>, a, sm_count, s);
        else launch_persist_k<LOGN>(k_ntt_persist<LOGN, false, ArithU64, false>, a, sm_count, s);
    }
}

template <int LOGN, class K>
static void launch_cluster_k(K kernel, const NttClArgs& a, cudaStream_t s) {
    typedef ClusterCfg<LOGN> C;
    constexpr int smem = C::TILE * 8;
    // (measured and dropped: padding the dynamic shared memory so that only 7 / 6 / 5 NTT CTAs fit an SM and the other
    //  stream's HBM-bound kernels of the two-stream pipeline co-reside: 6.04 / 6.31 / 6.63 ms per cfg2 step against 5.70;
    //  prefetch.global.L1 of the next round's twiddles before each barrier, 2464 vs 2601 GB/s --
    //  the kernel is short of issue slots, not of latency hiding; running every K-th FP64-eligible limb on the integer policy inside the same launch, to use
    //  both multiply pipes of an SM, was slower for every K -- 2.12 TB/s at K=4 against 2.60 TB/s; issue slots bind)
    LSB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (C::CL > 8) LSB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(a.total_rows * C::CL), 1, 1);
    cfg.blockDim = dim3(C::NT, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C::CL;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    static const bool report = [] {  // LSB_NTT_OCC=1: how many clusters of this kernel the device holds at once (stderr, once per kernel)
        const char* e = getenv("LSB_NTT_OCC");
        return e && atoi(e) != 0;
    }();
    if (report) {
        static std::mutex mu;
        static std::map<const void*, int> seen;
        std::lock_guard<std::mutex> lk(mu);
        if (!seen.count((const void*)kernel)) {
            int ncl = 0, per_sm = 0, dev = 0, sms = 0;
            cudaFuncAttributes fa;
            cudaOccupancyMaxActiveClusters(&ncl, kernel, &cfg);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, C::NT, smem);
            cudaFuncGetAttributes(&fa, kernel);
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            seen[(const void*)kernel] = ncl;
            fprintf(stderr, "lsb ntt occupancy: logN %d cluster %d x %d threads, %d KiB, %d registers: %d clusters = %.2f CTAs per SM co-resident "
                            "(%d per SM without the cluster constraint)\n", LOGN, C::CL, C::NT, smem >> 10, fa.numRegs, ncl, (double)ncl * C::CL / sms, per_sm);
        }
    }
    LSB_CUDA(cudaLaunchKernelEx(&cfg, kernel, a));
}

template <int LOGN>
static void launch_cluster_n(const NttClArgs& a, bool inverse, bool f64, cudaStream_t s) {
    if (a.npeer > 0) {
        if (f64) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithF64C, false, true>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithU64, false, true>, a, s);
        return;
    }
    if (a.fm) {
        if (f64) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithF64C, true>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithU64, true>, a, s);
        return;
    }
    if (f64) {
        if (inverse) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithF64C, false>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithF64C, false>, a, s);
    } else {
        if (inverse) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithU64, false>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithU64, false>, a, s);
    }
}

// Rows whose prime qualifies for the FP64 policy and the others go in separate launches (CTA-uniform code).
void launch_ntt_cluster(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch, bool inverse,
                        const std::vector<uint8_t>& f64_ok, bool allow_f64, cudaStream_t s0, const NttFinish* fin,
                        u64* const* peers, int npeer) {
    LSB_CHECK(!(fin && inverse), "lsb: the finish epilogue belongs to the forward transform");
    LSB_CHECK(npeer == 0 || (inverse && !fin && batch == 1 && npeer <= kMaxPeers && peers),
              "lsb: peer stores belong to the un-batched inverse transform (at most 8 peers)");
    RowMap parts[2];
    parts[0].n = parts[1].n = 0;
    for (int r = 0; r < map.n; r++) {
        RowMap& p = parts[(allow_f64 && f64_ok[map.mod[r]]) ? 1 : 0];
        p.in_slot[p.n] = map.in_slot[r];
        p.out_slot[p.n] = map.out_slot[r];
        p.mod[p.n] = map.mod[r];
        p.n++;
    }
    // A transform over limbs of both policies is two launches on disjoint rows: the FP64-policy one goes to a side stream so
    // that the two share the SMs (one is bound by the FP64 pipe, the other by the integer multiplier) and neither leaves
    // the machine half empty at its tail.  LSB_NTT_MIXED_OVERLAP=0: one after the other on `s0`.
    static const bool overlap = [] {
        const char* e = getenv("LSB_NTT_MIXED_OVERLAP");
        return !(e && atoi(e) == 0);
    }();
    const bool two = overlap && parts[0].n > 0 && parts[1].n > 0 && npeer == 0 && dt.ntt_side != nullptr;
    if (two) {
        LSB_CUDA(cudaEventRecord(dt.ntt_fork, s0));
        LSB_CUDA(cudaStreamWaitEvent(dt.ntt_side, dt.ntt_fork, 0));
    }
    for (int f = 0; f < 2; f++) {
        if (parts[f].n == 0) continue;
        cudaStream_t s = (two && f == 1) ? dt.ntt_side : s0;
        NttClArgs a;
        a.in = in;
        a.out = out;
        a.total_rows = batch * parts[f].n;
        a.tw = inverse ? dt.twc_inv : dt.twc_fwd;  // tail-SoA copies (host::soa_tail)
        a.tw_d = inverse ? dt.twc_inv_d : dt.twc_fwd_d;
        a.tw_w = inverse ? dt.twc_inv_w : dt.twc_fwd_w;
        a.mods = dt.mods;
        a.map = parts[f];
        a.fm = fin ? fin->fm_dev : nullptr;
        if (fin) {
            a.fa = fin->a;
            a.fadd = fin->addend;
            a.fout = fin->out;
            a.has_add = fin->has_addend ? 1 : 0;
        } else {
            a.fa = a.fadd = a.fout = in;
            a.has_add = 0;
        }
        a.npeer = npeer;
        for (int pr = 0; pr < npeer; pr++) a.peer[pr] = peers[pr];
        a.batch = batch;
        a.twp = inverse ? dt.tw_inv : dt.tw_fwd;
        a.twp_d = inverse ? dt.tw_inv_d : dt.tw_fwd_d;
        ProfScope _ps(dt.prof, inverse ? PK_NTT_INV : PK_NTT_FWD, s, 16.0 * dt.N * a.total_rows, 1);
        static const bool persist = [] {  // LSB_NTT_PERSIST=1: persistent clusters (k_ntt_persist) for N <= 2^14
            const char* e = getenv("LSB_NTT_PERSIST");
            return e && atoi(e) != 0;
        }();
        if (persist && npeer == 0 && dt.logN <= 14) {
            switch (dt.logN) {
                case 12: launch_persist_n<12>(a, inverse, f == 1, dt.sm_count, s); break;
                case 13: launch_persist_n<13>(a, inverse, f == 1, dt.sm_count, s); break;
                default: launch_persist_n<14>(a, inverse, f == 1, dt.sm_count, s); break;
            }
            continue;
        }
        switch (dt.logN) {
            case 12: launch_cluster_n<12>(a, inverse, f == 1, s); break;
            case 13: launch_cluster_n<13>(a, inverse, f == 1, s); break;
            case 14: launch_cluster_n<14>(a, inverse, f == 1, s); break;
            case 15: launch_cluster_n<15>(a, inverse, f == 1, s); break;
            case 16: launch_cluster_n<16>(a, inverse, f == 1, s); break;
            default: throw std::runtime_error("lsb: unsupported ring degree (logN must be 12..16)");
        }
    }
    if (two) {
        LSB_CUDA(cudaEventRecord(dt.ntt_join, dt.ntt_side));
        LSB_CUDA(cudaStreamWaitEvent(s0, dt.ntt_join, 0));
    }
}

}  // namespace lsb
