// Batched negacyclic NTT / INTT kernels for sm_100a (see lsb_ntt_core.h for the tile engine).
//
// Replaces the HEonGPU NTT the reference's GPU executor reaches through every operator
// (mega_ag_runners/gpu/mega_ag_executors_gpu.cu:185-289); results are limb-for-limb those of
// Lattigo's ring.NTT / ring.InvNTT as restated in oracle/orc_core.hpp:55-94.
//
//  * N <= 2^14: one CTA owns a whole limb in shared memory (<= 128 KiB): one HBM read + one HBM
//    write per limb (16*N bytes, the algorithmic minimum), persistent grid-stride loop over rows.
//  * N  = 2^15, 2^16: column pass (stages [0, logN-8) on 16-column strided tiles, 128-byte
//    segments) + row pass (last 8 stages on 4096-word contiguous tiles).
#include <atomic>
#include <type_traits>
#include <cstdlib>
#include "lsb_kernels.h"
#include "lsb_ntt_core.h"

namespace lsb {

template <int I, int NR, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (I < NR) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, NR>(f);
    }
}

struct NttArgs {
    BatchPtr in, out;
    int total_rows;
    const Tw* tw;  // [nmod][N]
    const ModConst* mods;
    RowMap map;
};

template <int LOGN>
struct SingleCfg {
    static constexpr int N = 1 << LOGN;
    // threads per CTA: groups are strided over threads, so NT only trades registers for warps
    static constexpr int NT = LOGN >= 14 ? 512 : 256;
};

template <int LOGN, bool INV>
__global__ void __launch_bounds__(SingleCfg<LOGN>::NT, 1) k_ntt_single(const __grid_constant__ NttArgs a) {
    extern __shared__ __align__(1024) u64 sm[];
    constexpr int N = 1 << LOGN;
    constexpr int NT = SingleCfg<LOGN>::NT;
    constexpr Plan P = make_plan(LOGN, 0);
    const int tid = threadIdx.x;
    const TileGeom tg{0, 0};
    for (int row = blockIdx.x; row < a.total_rows; row += gridDim.x) {
        const int b = row / a.map.n, r = row - b * a.map.n;
        const u64* src = a.in.at(b) + (size_t)a.map.in_slot[r] * N;
        u64* dst = a.out.at(b) + (size_t)a.map.out_slot[r] * N;
        const int mod = a.map.mod[r];
        const Tw* tw = a.tw + (size_t)mod * N;
        const u64 q = a.mods[mod].q;
        if (!INV) {
            const u64 qinv64 = a.mods[mod].qinv64;
            const bool lazy = q < kLazyLimit;  // CTA-uniform
            static_for<0, P.nrounds>([&](auto ic) {
                constexpr int i = decltype(ic)::value;
                constexpr int lr = P.rounds[i].lr;
                constexpr int loggc = plan_loggc(P, i);
                constexpr int SRC = i == 0 ? IO_GLOBAL : IO_SMEM;
                constexpr bool FINAL = i == P.nrounds - 1;
                if (lazy) {
                    for (int g = tid; g < (N >> lr); g += NT)
                        fwd_round<lr, SRC, IO_SMEM, FINAL, loggc == 0, true>(sm, src, nullptr, tg, tw, g, loggc,
                                                                            P.rounds[i].v, 1u, q, qinv64);
                } else {
                    for (int g = tid; g < (N >> lr); g += NT)
                        fwd_round<lr, SRC, IO_SMEM, FINAL, loggc == 0, false>(sm, src, nullptr, tg, tw, g, loggc,
                                                                             P.rounds[i].v, 1u, q);
                }
                __syncthreads();
            });
            for (int c = tid; c < N / 2; c += NT)
                reinterpret_cast<W2*>(dst)[c] = *reinterpret_cast<const W2*>(&sm[swz(2 * c)]);
            __syncthreads();
        } else {
            const Tw ninv = a.mods[mod].ninv, w_ninv = a.mods[mod].w_ninv;
            for (int c = tid; c < N / 2; c += NT)
                *reinterpret_cast<W2*>(&sm[swz(2 * c)]) = reinterpret_cast<const W2*>(src)[c];
            __syncthreads();
            static_for<0, P.nrounds>([&](auto ic) {
                constexpr int i = P.nrounds - 1 - decltype(ic)::value;
                constexpr int lr = P.rounds[i].lr;
                constexpr int loggc = plan_loggc(P, i);
                constexpr int DST = i == 0 ? IO_GLOBAL : IO_SMEM;
                for (int g = tid; g < (N >> lr); g += NT)
                    inv_round<lr, IO_SMEM, DST, i == 0, loggc == 0>(sm, nullptr, dst, tg, tw, g, loggc, P.rounds[i].v,
                                                                   1u, q, ninv, w_ninv);
                __syncthreads();
            });
        }
    }
}

// ---- two-pass (N = 2^15, 2^16) ----
constexpr int kRowLS = 8;    // row pass: 256-point sub-transforms
constexpr int kColLC = 4;    // column pass: 16 columns = 128-byte segments
constexpr int kRowSub = 16;  // sub-tiles per row-pass CTA (4096 words, 32 KiB)

template <int LOGN, bool INV>
__global__ void __launch_bounds__((1 << (LOGN - kRowLS + kColLC)) / 16) k_ntt_cols(const __grid_constant__ NttArgs a) {
    extern __shared__ __align__(1024) u64 sm[];
    constexpr int N = 1 << LOGN;
    constexpr int LA = LOGN - kRowLS;
    constexpr int T = 1 << (LA + kColLC);
    constexpr int NT = T / 16;
    constexpr int UNITS = (1 << kRowLS) >> kColLC;  // column blocks per row
    constexpr Plan P = make_plan(LA, kColLC);
    const int tid = threadIdx.x;
    const TileGeom tg{kColLC, kRowLS};
    for (int unit = blockIdx.x; unit < a.total_rows * UNITS; unit += gridDim.x) {
        const int row = unit / UNITS, cb = unit - row * UNITS;
        const int b = row / a.map.n, r = row - b * a.map.n;
        const u64* src = a.in.at(b) + (size_t)a.map.in_slot[r] * N + (cb << kColLC);
        u64* dst = a.out.at(b) + (size_t)a.map.out_slot[r] * N + (cb << kColLC);
        const int mod = a.map.mod[r];
        const Tw* tw = a.tw + (size_t)mod * N;
        const u64 q = a.mods[mod].q;
        const Tw ninv = a.mods[mod].ninv, w_ninv = a.mods[mod].w_ninv;
        static_for<0, P.nrounds>([&](auto ic) {
            constexpr int j = decltype(ic)::value;
            constexpr int i = INV ? P.nrounds - 1 - j : j;
            constexpr int lr = P.rounds[i].lr;
            constexpr int loggc = plan_loggc(P, i);
            constexpr int SRC = j == 0 ? IO_GLOBAL : IO_SMEM;
            constexpr int DST = j == P.nrounds - 1 ? IO_GLOBAL : IO_SMEM;
            for (int g = tid; g < (T >> lr); g += NT) {
                if (!INV) {
                    if (q < kLazyLimit)
                        fwd_round<lr, SRC, DST, false, false, true>(sm, src, dst, tg, tw, g, loggc, P.rounds[i].v, 1u, q);
                    else
                        fwd_round<lr, SRC, DST, false, false, false>(sm, src, dst, tg, tw, g, loggc, P.rounds[i].v, 1u, q);
                } else
                    inv_round<lr, SRC, DST, i == 0, false>(sm, src, dst, tg, tw, g, loggc, P.rounds[i].v, 1u, q, ninv,
                                                           w_ninv);
            }
            __syncthreads();
        });
    }
}

template <int LOGN, bool INV>
__global__ void __launch_bounds__(256) k_ntt_rows(const __grid_constant__ NttArgs a) {
    extern __shared__ __align__(1024) u64 sm[];
    constexpr int N = 1 << LOGN;
    constexpr int LA = LOGN - kRowLS;
    constexpr int S = 1 << kRowLS;
    constexpr int T = S * kRowSub;
    constexpr int NT = 256;
    constexpr int UNITS = N / T;
    constexpr Plan P = make_plan(kRowLS, 0);  // [4,4]
    static_assert(P.nrounds == 2 && P.rounds[0].lr == 4 && P.rounds[1].lr == 4, "row pass expects two radix-16 rounds");
    const int tid = threadIdx.x;
    const int st = tid >> 4, g = tid & 15;
    const TileGeom tg{0, 0};
    for (int unit = blockIdx.x; unit < a.total_rows * UNITS; unit += gridDim.x) {
        const int row = unit / UNITS, tb = unit - row * UNITS;
        const int b = row / a.map.n, r = row - b * a.map.n;
        const u64* src = a.in.at(b) + (size_t)a.map.in_slot[r] * N + (size_t)tb * T;
        u64* dst = a.out.at(b) + (size_t)a.map.out_slot[r] * N + (size_t)tb * T;
        const int mod = a.map.mod[r];
        const Tw* tw = a.tw + (size_t)mod * N;
        const u64 q = a.mods[mod].q;
        const u32 twbase = (1u << LA) + (u32)(tb * kRowSub + st);
        u64* smt = sm + st * S;
        if (!INV) {
            const u64 qinv64 = a.mods[mod].qinv64;
            if (q < kLazyLimit) {
                fwd_round<4, IO_GLOBAL, IO_SMEM, false, false, true>(smt, src + st * S, nullptr, tg, tw, g, 4, 0, twbase, q);
                __syncthreads();
                fwd_round<4, IO_SMEM, IO_SMEM, true, true, true>(smt, nullptr, nullptr, tg, tw, g, 0, 4, twbase, q, qinv64);
            } else {
                fwd_round<4, IO_GLOBAL, IO_SMEM, false, false, false>(smt, src + st * S, nullptr, tg, tw, g, 4, 0, twbase, q);
                __syncthreads();
                fwd_round<4, IO_SMEM, IO_SMEM, true, true, false>(smt, nullptr, nullptr, tg, tw, g, 0, 4, twbase, q);
            }
            __syncthreads();
            for (int c = tid; c < T / 2; c += NT)
                reinterpret_cast<W2*>(dst)[c] = *reinterpret_cast<const W2*>(&sm[swz(2 * c)]);
            __syncthreads();
        } else {
            const Tw z{0, 0};
            for (int c = tid; c < T / 2; c += NT)
                *reinterpret_cast<W2*>(&sm[swz(2 * c)]) = reinterpret_cast<const W2*>(src)[c];
            __syncthreads();
            inv_round<4, IO_SMEM, IO_SMEM, false, true>(smt, nullptr, nullptr, tg, tw, g, 0, 4, twbase, q, z, z);
            __syncthreads();
            inv_round<4, IO_SMEM, IO_GLOBAL, false, false>(smt, nullptr, dst + st * S, tg, tw, g, 4, 0, twbase, q, z, z);
            __syncthreads();
        }
    }
}

constexpr int kMaxDevices = 64;
static int current_device() {
    int dev = 0;
    LSB_CUDA(cudaGetDevice(&dev));
    LSB_CHECK(dev >= 0 && dev < kMaxDevices, "lsb: device ordinal out of range");
    return dev;
}

// ---------------------------------------------------------------------------------------------
template <int LOGN, bool INV>
static void launch_single(const NttArgs& a, int sms, cudaStream_t s) {
    constexpr int smem = (1 << LOGN) * 8;
    static std::atomic<bool> attr_done[kMaxDevices];  // function attributes are per device
    const int dev = current_device();
    if (!attr_done[dev].load(std::memory_order_acquire)) {
        LSB_CUDA(cudaFuncSetAttribute(k_ntt_single<LOGN, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done[dev].store(true, std::memory_order_release);
    }
    int per_sm = (200 * 1024) / smem;
    if (per_sm < 1) per_sm = 1;
    int grid = a.total_rows < sms * per_sm ? a.total_rows : sms * per_sm;
    k_ntt_single<LOGN, INV><<<grid, SingleCfg<LOGN>::NT, smem, s>>>(a);
    LSB_CUDA(cudaGetLastError());
}

template <int LOGN, bool INV>
static void launch_two_pass(NttArgs a, int sms, cudaStream_t s) {
    constexpr int LA = LOGN - kRowLS;
    constexpr int smem_c = (1 << (LA + kColLC)) * 8, nt_c = (1 << (LA + kColLC)) / 16;
    constexpr int smem_r = (1 << kRowLS) * kRowSub * 8;
    static std::atomic<bool> attr_done[kMaxDevices];  // function attributes are per device
    const int dev = current_device();
    if (!attr_done[dev].load(std::memory_order_acquire)) {
        LSB_CUDA(cudaFuncSetAttribute(k_ntt_cols<LOGN, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_c));
        LSB_CUDA(cudaFuncSetAttribute(k_ntt_rows<LOGN, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_r));
        attr_done[dev].store(true, std::memory_order_release);
    }
    const int units_c = a.total_rows * ((1 << kRowLS) >> kColLC);
    const int units_r = a.total_rows * ((1 << LOGN) / ((1 << kRowLS) * kRowSub));
    const int gc = units_c < sms * 8 ? units_c : sms * 8;
    const int gr = units_r < sms * 8 ? units_r : sms * 8;
    if (!INV) {
        k_ntt_cols<LOGN, false><<<gc, nt_c, smem_c, s>>>(a);
        LSB_CUDA(cudaGetLastError());
        NttArgs a2 = a;  // second pass works in place on the output
        a2.in = a.out;
        for (int i = 0; i < a.map.n; i++) a2.map.in_slot[i] = a.map.out_slot[i];
        k_ntt_rows<LOGN, false><<<gr, 256, smem_r, s>>>(a2);
    } else {
        k_ntt_rows<LOGN, true><<<gr, 256, smem_r, s>>>(a);
        LSB_CUDA(cudaGetLastError());
        NttArgs a2 = a;
        a2.in = a.out;
        for (int i = 0; i < a.map.n; i++) a2.map.in_slot[i] = a.map.out_slot[i];
        k_ntt_cols<LOGN, true><<<gc, nt_c, smem_c, s>>>(a2);
    }
    LSB_CUDA(cudaGetLastError());
}

static bool ntt_allow_f64() {
    static const bool v = [] {
        const char* e = getenv("LSB_NTT_F64");
        return !(e && atoi(e) == 0);
    }();
    return v;
}
static bool ntt_use_cluster() {
    static const bool v = [] {
        const char* e = getenv("LSB_NTT_IMPL");
        return !(e && std::string(e) == "single");
    }();
    return v;
}

// forward NTT of b[fm.b_slot[r]] + finish; with the cluster kernel the finish is its epilogue (LSB_NO_EPILOGUE=1: two launches)
void launch_ntt_finish(const DeviceTables& dt, BatchPtr b, const NttFinish& fin, int batch, cudaStream_t s) {
    const FinishMap& fm = *fin.fm_host;
    if (batch <= 0 || fm.n <= 0) return;
    static const bool fused = [] {
        const char* e = getenv("LSB_NO_EPILOGUE");
        return !(e && atoi(e) != 0);
    }();
    RowMap m;
    m.n = fm.n;
    for (int r = 0; r < fm.n; r++) {
        m.in_slot[r] = fm.b_slot[r];
        m.out_slot[r] = fused && ntt_use_cluster() ? (uint8_t)r : fm.b_slot[r];
        m.mod[r] = fm.mod[r];
    }
    if (fused && ntt_use_cluster()) {
        launch_ntt_cluster(dt, b, b, m, batch, false, dt.f64_ok, ntt_allow_f64(), s, &fin);
        return;
    }
    launch_ntt(dt, b, b, m, batch, false, s);
    launch_finish(dt, fin.a, b, fin.addend, fin.has_addend, fin.out, fin.fm_dev, fm.n, batch, s);
}

void launch_intt_to_peers(const DeviceTables& dt, BatchPtr in, const RowMap& map, u64* const* peers, int npeer,
                          cudaStream_t s) {
    if (map.n <= 0) return;
    launch_ntt_cluster(dt, in, in, map, 1, true, dt.f64_ok, ntt_allow_f64(), s, nullptr, peers, npeer);
}

static void launch_ntt_first_round(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch,
                                   bool inverse, cudaStream_t s);

void launch_ntt(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch, bool inverse,
                cudaStream_t s) {
    if (batch <= 0 || map.n <= 0) return;
    static const int impl = [] {  // 0: cluster kernel (default), 1: single-CTA / two-pass kernels
        const char* e = getenv("LSB_NTT_IMPL");
        return (e && std::string(e) == "single") ? 1 : 0;
    }();
    static const bool allow_f64 = [] {
        const char* e = getenv("LSB_NTT_F64");
        return !(e && atoi(e) == 0);
    }();
    if (impl == 0) {
        // Integer-policy rows of a forward transform at N = 2^14 / 2^16 are faster on the first-round kernels
        // (measured, 61-bit primes: 1355 vs 1201 GB/s at 2^14, 1047 vs 847 GB/s at 2^16); everything else -- all
        // FP64-eligible rows, all inverses, N = 2^12, 2^13, 2^15 -- runs on the cluster kernel.
        static const bool route = [] {
            const char* e = getenv("LSB_NTT_ROUTE");
            return !(e && atoi(e) == 0);
        }();
        if (route && !inverse && (dt.logN == 14 || dt.logN == 16)) {
            RowMap cl, sg;
            cl.n = sg.n = 0;
            for (int r = 0; r < map.n; r++) {
                RowMap& p = (allow_f64 && dt.f64_ok[map.mod[r]]) ? cl : sg;
                p.in_slot[p.n] = map.in_slot[r];
                p.out_slot[p.n] = map.out_slot[r];
                p.mod[p.n] = map.mod[r];
                p.n++;
            }
            // both kinds of rows: the FP64-policy launch goes to the side stream beside the integer-policy one (as in
            // launch_ntt_cluster; LSB_NTT_MIXED_OVERLAP=0 keeps them one after the other)
            static const bool overlap = [] {
                const char* e = getenv("LSB_NTT_MIXED_OVERLAP");
                return !(e && atoi(e) == 0);
            }();
            const bool two = overlap && cl.n > 0 && sg.n > 0 && dt.ntt_side != nullptr;
            if (two) {
                LSB_CUDA(cudaEventRecord(dt.ntt_fork, s));
                LSB_CUDA(cudaStreamWaitEvent(dt.ntt_side, dt.ntt_fork, 0));
            }
            if (cl.n) launch_ntt_cluster(dt, in, out, cl, batch, false, dt.f64_ok, allow_f64, two ? dt.ntt_side : s);
            if (sg.n) launch_ntt_first_round(dt, in, out, sg, batch, false, s);
            if (two) {
                LSB_CUDA(cudaEventRecord(dt.ntt_join, dt.ntt_side));
                LSB_CUDA(cudaStreamWaitEvent(s, dt.ntt_join, 0));
            }
            return;
        }
        launch_ntt_cluster(dt, in, out, map, batch, inverse, dt.f64_ok, allow_f64, s);
        return;
    }
    launch_ntt_first_round(dt, in, out, map, batch, inverse, s);
}

static void launch_ntt_first_round(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch,
                                   bool inverse, cudaStream_t s) {
    NttArgs a;
    a.in = in;
    a.out = out;
    a.total_rows = batch * map.n;
    a.tw = inverse ? dt.tw_inv : dt.tw_fwd;
    a.mods = dt.mods;
    a.map = map;
    const int sms = dt.sm_count;
    ProfScope _ps(dt.prof, inverse ? PK_NTT_INV : PK_NTT_FWD, s, 16.0 * dt.N * a.total_rows, dt.logN > 14 ? 2 : 1);
#define LSB_NTT_CASE1(L)                                  \
    case L:                                               \
        if (inverse) launch_single<L, true>(a, sms, s);   \
        else launch_single<L, false>(a, sms, s);          \
        break;
#define LSB_NTT_CASE2(L)                                  \
    case L:                                               \
        if (inverse) launch_two_pass<L, true>(a, sms, s); \
        else launch_two_pass<L, false>(a, sms, s);        \
        break;
    static const bool two14 = [] {
        const char* e = getenv("LSB_NTT_TWOPASS");
        return e && atoi(e) != 0;
    }();
    if (dt.logN == 14 && two14) {
        if (inverse) launch_two_pass<14, true>(a, sms, s);
        else launch_two_pass<14, false>(a, sms, s);
        return;
    }
    switch (dt.logN) {
        LSB_NTT_CASE1(12)
        LSB_NTT_CASE1(13)
        LSB_NTT_CASE1(14)
        LSB_NTT_CASE2(15)
        LSB_NTT_CASE2(16)
        default: throw std::runtime_error("lsb: unsupported ring degree (logN must be 12..16)");
    }
}

}  // namespace lsb
