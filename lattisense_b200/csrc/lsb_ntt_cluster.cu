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
#include <type_traits>

#include "lsb_kernels.h"
#include "lsb_ntt_core.h"
#include "lsb_ntt_cluster_core.h"

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
namespace cg = cooperative_groups;

namespace lsb {

template <int LOGN>
struct ClusterCfg {
    static constexpr int N = 1 << LOGN;
    static constexpr int CL = LOGN <= 12 ? 1 : LOGN == 13 ? 2 : LOGN == 14 ? LSB_CL14 : 8;  // CTAs per limb
    static constexpr int BS = N / 16;                                                 // block = sub-transform size
    static constexpr int LS = LOGN - 4;
    static constexpr int NB = 16 / CL;      // blocks per CTA
    static constexpr int TILE = N / CL;     // words of shared memory per CTA
    static constexpr int NT = LOGN == 14 ? LSB_CL14_NT : LSB_CL_NT;
    // resident CTAs per SM asked of ptxas; N = 2^16 tiles are 64 KiB, so only 3 fit anyway and asking for 3 leaves the
    // compiler 85 registers (inverse: 1365 -> 1538 GB/s on the mixed chain, 1004 -> 1098 GB/s on 61-bit primes)
    static constexpr int MINB = LOGN == 14 ? LSB_CL14_MINB : LOGN == 16 ? 3 : LSB_CL_MINB;
    static constexpr int G0 = BS / CL;      // cross-round groups per CTA
};

constexpr int kMaxPeers = 8;
struct NttClArgs {
    BatchPtr in, out;
    int total_rows;
    const Tw* tw;     // [nmod][N]
    const TwD* tw_d;  // [nmod][N]
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
    if constexpr (std::is_same<P, ArithF64>::value) return a.tw_d + (size_t)mod * N;
    else return a.tw + (size_t)mod * N;
}

template <int LOGN, bool INV, class P, bool FIN, bool PEERS = false>
__device__ __forceinline__ void ntt_cluster_body(const NttClArgs& a, u64* sm) {
    typedef ClusterCfg<LOGN> C;
    constexpr int N = C::N, CL = C::CL, BS = C::BS, NB = C::NB, TILE = C::TILE, NT = C::NT;
    constexpr Plan PL = make_plan(C::LS, 0);
    const int tid = threadIdx.x;
    const int rank = CL > 1 ? (int)(blockIdx.x % CL) : 0;
    const int row = blockIdx.x / CL;
    const int b = row / a.map.n, r = row - b * a.map.n;
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
        if constexpr (std::is_same<P, ArithF64>::value) {
            ninv = mc.ninv_d;
            w_ninv = mc.w_ninv_d;
        } else {
            ninv = mc.ninv;
            w_ninv = mc.w_ninv;
        }
        for (int gl = tid; gl < C::G0; gl += NT) {
            const int g = rank * C::G0 + gl;
            typename P::V x[16];
#pragma unroll
            for (int k = 0; k < 16; k++) x[k] = P::from_bits(peer[k / NB][(k % NB) * BS + swz(g)]);
            inv_stages<P, 4, true>(x, tw, 1u, m, ninv, w_ninv);
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
        if (CL > 1) cluster.sync();  // peers may still be reading this CTA's shared memory
    }
}

template <int LOGN, bool INV, class P, bool FIN = false, bool PEERS = false>
__global__ void __launch_bounds__(ClusterCfg<LOGN>::NT, ClusterCfg<LOGN>::MINB) k_ntt_cluster(const __grid_constant__ NttClArgs a) {
    extern __shared__ __align__(1024) u64 sm[];
    ntt_cluster_body<LOGN, INV, P, FIN, PEERS>(a, sm);
}

template <int LOGN, class K>
static void launch_cluster_k(K kernel, const NttClArgs& a, cudaStream_t s) {
    typedef ClusterCfg<LOGN> C;
    constexpr int smem = C::TILE * 8;
    // (measured and dropped: prefetch.global.L1 of the next round's twiddles before each barrier, 2464 vs 2601 GB/s --
    //  the kernel is short of issue slots, not of latency hiding; running every K-th FP64-eligible limb on the integer policy inside the same launch, to use
    //  both multiply pipes of an SM, was slower for every K -- 2.12 TB/s at K=4 against 2.60 TB/s; issue slots bind)
    LSB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
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
    LSB_CUDA(cudaLaunchKernelEx(&cfg, kernel, a));
}

template <int LOGN>
static void launch_cluster_n(const NttClArgs& a, bool inverse, bool f64, cudaStream_t s) {
    if (a.npeer > 0) {
        if (f64) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithF64, false, true>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithU64, false, true>, a, s);
        return;
    }
    if (a.fm) {
        if (f64) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithF64, true>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithU64, true>, a, s);
        return;
    }
    if (f64) {
        if (inverse) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithF64, false>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithF64, false>, a, s);
    } else {
        if (inverse) launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, true, ArithU64, false>, a, s);
        else launch_cluster_k<LOGN>(k_ntt_cluster<LOGN, false, ArithU64, false>, a, s);
    }
}

// Rows whose prime qualifies for the FP64 policy and the others go in separate launches (CTA-uniform code).
void launch_ntt_cluster(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch, bool inverse,
                        const std::vector<uint8_t>& f64_ok, bool allow_f64, cudaStream_t s, const NttFinish* fin,
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
    for (int f = 0; f < 2; f++) {
        if (parts[f].n == 0) continue;
        NttClArgs a;
        a.in = in;
        a.out = out;
        a.total_rows = batch * parts[f].n;
        a.tw = inverse ? dt.twc_inv : dt.twc_fwd;  // tail-SoA copies (host::soa_tail)
        a.tw_d = inverse ? dt.twc_inv_d : dt.twc_fwd_d;
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
        ProfScope _ps(dt.prof, inverse ? PK_NTT_INV : PK_NTT_FWD, s, 16.0 * dt.N * a.total_rows, 1);
        switch (dt.logN) {
            case 12: launch_cluster_n<12>(a, inverse, f == 1, s); break;
            case 13: launch_cluster_n<13>(a, inverse, f == 1, s); break;
            case 14: launch_cluster_n<14>(a, inverse, f == 1, s); break;
            case 15: launch_cluster_n<15>(a, inverse, f == 1, s); break;
            case 16: launch_cluster_n<16>(a, inverse, f == 1, s); break;
            default: throw std::runtime_error("lsb: unsupported ring degree (logN must be 12..16)");
        }
    }
}

}  // namespace lsb
