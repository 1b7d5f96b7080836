// Limb-wise and per-coefficient kernels around the NTT: tensor, add/sub/neg, plaintext multiply,
// basis extension (ModUp / ModDown), key-switch inner product, rescale, automorphisms.
// All are HBM-streaming kernels: 128-bit coalesced accesses, grid = (coefficient blocks, rows, batch).
//
// Reference call sites replaced: mega_ag_runners/gpu/mega_ag_executors_gpu.cu:71-175 (add/sub/neg),
// :177-227 (multiply), :229-238 (relinearize), :240-248 (rescale), :264-292 (rotations).
#include "lsb_elem_core.h"
#include "lsb_ntt_core.h"  // W2
#include "lsb_ntt_arith.h"

namespace lsb {

constexpr int kElemThreads = 256;

static bool elem_f64_enabled() {  // LSB_ELEM_F64=0: integer arithmetic in every element-wise kernel (A/B switch)
    static const bool v = [] {
        const char* e = getenv("LSB_ELEM_F64");
        return !(e && atoi(e) == 0);
    }();
    return v;
}

static inline dim3 elem_grid(int N, int rows, int batch) { return dim3((N / 2 + kElemThreads - 1) / kElemThreads, rows, batch); }

__device__ __forceinline__ W2 ld2(const u64* p, int j2) { return reinterpret_cast<const W2*>(p)[j2]; }
__device__ __forceinline__ void st2(u64* p, int j2, W2 v) { reinterpret_cast<W2*>(p)[j2] = v; }

// ---- add / sub / neg -------------------------------------------------------------------------
struct BinArgs {
    BatchPtr a, b, out;
    int N, b_rows, op;
    const ModConst* mods;
    RowMap map;
};
__global__ void __launch_bounds__(kElemThreads) k_binary(const __grid_constant__ BinArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int r = blockIdx.y, b = blockIdx.z;
    const u64 q = g.mods[g.map.mod[r]].q;
    const u64* pa = g.a.at(b) + (size_t)g.map.in_slot[r] * g.N;
    u64* po = g.out.at(b) + (size_t)g.map.out_slot[r] * g.N;
    W2 x = ld2(pa, j2);
    if (r < g.b_rows) {
        W2 y = ld2(g.b.at(b) + (size_t)g.map.in_slot[r] * g.N, j2);
        if (g.op == OP_ADD) {
            x.a = addmod(x.a, y.a, q);
            x.b = addmod(x.b, y.b, q);
        } else {
            x.a = submod(x.a, y.a, q);
            x.b = submod(x.b, y.b, q);
        }
    }
    st2(po, j2, x);
}
void launch_binary(const DeviceTables& dt, BinOp op, BatchPtr a, BatchPtr b, BatchPtr out, const RowMap& map,
                   int b_rows, int batch, cudaStream_t s) {
    if (batch <= 0 || map.n <= 0) return;
    BinArgs g{a, b, out, dt.N, b_rows, (int)op, dt.mods, map};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * map.n * batch * 3);
    k_binary<<<elem_grid(dt.N, map.n, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

struct NegArgs {
    BatchPtr a, out;
    int N;
    const ModConst* mods;
    RowMap map;
};
__global__ void __launch_bounds__(kElemThreads) k_negate(const __grid_constant__ NegArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int r = blockIdx.y, b = blockIdx.z;
    const u64 q = g.mods[g.map.mod[r]].q;
    W2 x = ld2(g.a.at(b) + (size_t)g.map.in_slot[r] * g.N, j2);
    x.a = negmod(x.a, q);
    x.b = negmod(x.b, q);
    st2(g.out.at(b) + (size_t)g.map.out_slot[r] * g.N, j2, x);
}
void launch_negate(const DeviceTables& dt, BatchPtr a, BatchPtr out, const RowMap& map, int batch, cudaStream_t s) {
    if (batch <= 0 || map.n <= 0) return;
    NegArgs g{a, out, dt.N, dt.mods, map};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * map.n * batch * 2);
    k_negate<<<elem_grid(dt.N, map.n, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- tensor ------------------------------------------------------------------------------------
struct TensorArgs {
    BatchPtr a, b, out;
    int N, nl, in_ps, out_ps;
    const ModConst* mods;
    uint8_t mod[64];
};
template <bool F64>
__global__ void __launch_bounds__(kElemThreads) k_tensor(const __grid_constant__ TensorArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int i = blockIdx.y, b = blockIdx.z;
    const ModConst& mc = g.mods[g.mod[i]];
    const BarrettQ bq = mc.bq;
    const u64 q = bq.q;
    const u64* pa = g.a.at(b);
    const u64* pb = g.b.at(b);
    u64* po = g.out.at(b);
    const size_t N = g.N;
    const W2 a0 = ld2(pa + (size_t)i * N, j2), a1 = ld2(pa + (size_t)(g.in_ps + i) * N, j2);
    const W2 b0 = ld2(pb + (size_t)i * N, j2), b1 = ld2(pb + (size_t)(g.in_ps + i) * N, j2);
    W2 c0, c1, c2;
    if constexpr (F64) {
        // primes below 2^47: the four products per coefficient on the FP64 pipe (the integer version is bound by IMAD,
        // profiles/r01_step_kernels_ncu.txt); |a0 b1 + a1 b0| <= 3 q before the canonical reduction
        const double qd = mc.qd, qi = mc.qinvd;
        const double x0a = f64_from_u64(a0.a), x0b = f64_from_u64(a0.b), x1a = f64_from_u64(a1.a), x1b = f64_from_u64(a1.b);
        const double y0a = f64_from_u64(b0.a), y0b = f64_from_u64(b0.b), y1a = f64_from_u64(b1.a), y1b = f64_from_u64(b1.b);
        c0.a = f64_canon_u(f64_mulmod_var(x0a, y0a, qd, qi), qd, qi);
        c0.b = f64_canon_u(f64_mulmod_var(x0b, y0b, qd, qi), qd, qi);
        c2.a = f64_canon_u(f64_mulmod_var(x1a, y1a, qd, qi), qd, qi);
        c2.b = f64_canon_u(f64_mulmod_var(x1b, y1b, qd, qi), qd, qi);
        c1.a = f64_canon_u(d_add(f64_mulmod_var(x0a, y1a, qd, qi), f64_mulmod_var(x1a, y0a, qd, qi)), qd, qi);
        c1.b = f64_canon_u(d_add(f64_mulmod_var(x0b, y1b, qd, qi), f64_mulmod_var(x1b, y0b, qd, qi)), qd, qi);
        st2(po + (size_t)i * N, j2, c0);
        st2(po + (size_t)(g.out_ps + i) * N, j2, c1);
        st2(po + (size_t)(2 * g.out_ps + i) * N, j2, c2);
        return;
    }
    c0.a = mulmod(a0.a, b0.a, bq);
    c0.b = mulmod(a0.b, b0.b, bq);
    c2.a = mulmod(a1.a, b1.a, bq);
    c2.b = mulmod(a1.b, b1.b, bq);
    {
        Acc128 t{0, 0};
        acc_mad(t, a0.a, b1.a);
        acc_mad(t, a1.a, b0.a);
        c1.a = reduce_wide(t.lo, t.hi, bq);  // 2 q^2 < 2^(sh+64) for q < 2^62
        Acc128 u{0, 0};
        acc_mad(u, a0.b, b1.b);
        acc_mad(u, a1.b, b0.b);
        c1.b = reduce_wide(u.lo, u.hi, bq);
    }
    (void)q;
    st2(po + (size_t)i * N, j2, c0);
    st2(po + (size_t)(g.out_ps + i) * N, j2, c1);
    st2(po + (size_t)(2 * g.out_ps + i) * N, j2, c2);
}
void launch_tensor(const DeviceTables& dt, BatchPtr a, BatchPtr b, BatchPtr out, int nl, int a_poly_stride,
                   int out_poly_stride, const uint8_t* mod_of_limb, int batch, cudaStream_t s) {
    if (batch <= 0) return;
    LSB_CHECK(nl <= 64, "tensor: too many limbs");
    TensorArgs g;
    g.a = a;
    g.b = b;
    g.out = out;
    g.N = dt.N;
    g.nl = nl;
    g.in_ps = a_poly_stride;
    g.out_ps = out_poly_stride;
    g.mods = dt.mods;
    bool f64 = elem_f64_enabled();
    for (int i = 0; i < nl; i++) {
        g.mod[i] = mod_of_limb[i];
        f64 = f64 && dt.elem_f64_ok[mod_of_limb[i]];
    }
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * nl * batch * 7);
    if (f64) k_tensor<true><<<elem_grid(dt.N, nl, batch), kElemThreads, 0, s>>>(g);
    else k_tensor<false><<<elem_grid(dt.N, nl, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- ct x pt -----------------------------------------------------------------------------------
struct MulPlainArgs {
    BatchPtr ct, pt, out;
    int N, npoly, nl, acc;
    const ModConst* mods;
};
__global__ void __launch_bounds__(kElemThreads) k_mul_plain(const __grid_constant__ MulPlainArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int i = blockIdx.y, b = blockIdx.z;
    const BarrettQ bq = g.mods[i].bq;
    const W2 p = ld2(g.pt.at(b) + (size_t)i * g.N, j2);
    for (int pl = 0; pl < g.npoly; pl++) {
        const size_t off = (size_t)(pl * g.nl + i) * g.N;
        const W2 c = ld2(g.ct.at(b) + off, j2);
        W2 o;
        o.a = mulmod(c.a, p.a, bq);
        o.b = mulmod(c.b, p.b, bq);
        if (g.acc) {
            const W2 prev = ld2(g.out.at(b) + off, j2);
            o.a = addmod(o.a, prev.a, bq.q);
            o.b = addmod(o.b, prev.b, bq.q);
        }
        st2(g.out.at(b) + off, j2, o);
    }
}
void launch_mul_plain(const DeviceTables& dt, BatchPtr ct, BatchPtr pt, BatchPtr out, int npoly, int nl,
                      bool accumulate, int batch, cudaStream_t s) {
    if (batch <= 0) return;
    MulPlainArgs g{ct, pt, out, dt.N, npoly, nl, accumulate ? 1 : 0, dt.mods};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * nl * batch * (1 + 2 * npoly));
    k_mul_plain<<<elem_grid(dt.N, nl, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- n-ary sum: out = sum_t ct_t * pt_t + sum_u ct_u  (one pass over every term) --------------------
// The runner re-associates chains of ciphertext additions (and the ct x pt products feeding them) into one of
// these: modular addition is associative, so the limbs equal those of the node-by-node evaluation.
struct SumTermsArgs {
    const u64* const* ct;  // [n_mac + n_plain][batch]
    const u64* const* pt;  // [n_mac][batch]
    u64* const* out;       // [batch]
    int N, nl, n_mac, n_plain, batch;
    const ModConst* mods;
    // split > 1: the terms are cut into `split` ranges (chunk_mac / chunk_plain terms each), blockIdx.z = b * split + s
    // writes its partial sum to partial[(b * split + s) * words]; k_sum_partials adds them up.  Gives a long sum of one
    // ciphertext (12 800 terms in the convolution benchmark) enough CTAs to fill the machine.
    int split, chunk_mac, chunk_plain;
    u64* partial;
    size_t words;
};
__global__ void __launch_bounds__(kElemThreads) k_sum_terms(const __grid_constant__ SumTermsArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int r = blockIdx.y, z = blockIdx.z, b = z / g.split, sp = z - b * g.split, i = r % g.nl;
    const BarrettQ bq = g.mods[i].bq;
    const size_t coff = (size_t)r * g.N, poff = (size_t)i * g.N;
    const int m0 = min(g.n_mac, sp * g.chunk_mac), m1 = min(g.n_mac, m0 + g.chunk_mac);
    const int p0 = min(g.n_plain, sp * g.chunk_plain), p1 = min(g.n_plain, p0 + g.chunk_plain);
    W2 acc{0, 0};
#pragma unroll 4
    for (int t = m0; t < m1; t++) {
        const W2 c = ld2(g.ct[(size_t)t * g.batch + b] + coff, j2);
        const W2 p = ld2(g.pt[(size_t)t * g.batch + b] + poff, j2);
        acc.a = addmod(acc.a, mulmod(c.a, p.a, bq), bq.q);
        acc.b = addmod(acc.b, mulmod(c.b, p.b, bq), bq.q);
    }
#pragma unroll 4
    for (int t = p0; t < p1; t++) {
        const W2 c = ld2(g.ct[(size_t)(g.n_mac + t) * g.batch + b] + coff, j2);
        acc.a = addmod(acc.a, c.a, bq.q);
        acc.b = addmod(acc.b, c.b, bq.q);
    }
    u64* dst = g.split > 1 ? g.partial + (size_t)z * g.words : g.out[b];
    st2(dst + coff, j2, acc);
}
__global__ void __launch_bounds__(kElemThreads) k_sum_partials(const __grid_constant__ SumTermsArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int r = blockIdx.y, b = blockIdx.z;
    const u64 q = g.mods[r % g.nl].q;
    const size_t coff = (size_t)r * g.N;
    W2 acc{0, 0};
    for (int sp = 0; sp < g.split; sp++) {
        const W2 c = ld2(g.partial + ((size_t)b * g.split + sp) * g.words + coff, j2);
        acc.a = addmod(acc.a, c.a, q);
        acc.b = addmod(acc.b, c.b, q);
    }
    st2(g.out[b] + coff, j2, acc);
}
int sum_terms_split(const DeviceTables& dt, int npoly, int nl, int n_terms, int batch) {
    const long ctas = (long)((dt.N / 2 + kElemThreads - 1) / kElemThreads) * npoly * nl * batch;
    if (n_terms < 64 || ctas >= 4096) return 1;
    return (int)std::min<long>((n_terms + 31) / 32, (8192 + ctas - 1) / ctas);
}
void launch_sum_terms(const DeviceTables& dt, const u64* const* ct, const u64* const* pt, u64* const* out, int npoly, int nl,
                      int n_mac, int n_plain, int batch, cudaStream_t s, int split, u64* partial) {
    if (batch <= 0) return;
    if (split < 1 || !partial) split = 1;
    SumTermsArgs g{ct, pt, out, dt.N, nl, n_mac, n_plain, batch, dt.mods, split, (n_mac + split - 1) / split, (n_plain + split - 1) / split,
                   partial, (size_t)npoly * nl * dt.N};
    if (split == 1) g.chunk_mac = n_mac, g.chunk_plain = n_plain;
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * nl * batch * (npoly * (n_mac + n_plain + 1.0) + n_mac));
    k_sum_terms<<<elem_grid(dt.N, npoly * nl, batch * split), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
    if (split > 1) {
        k_sum_partials<<<elem_grid(dt.N, npoly * nl, batch), kElemThreads, 0, s>>>(g);
        LSB_CUDA(cudaGetLastError());
    }
}

// ---- basis extension -----------------------------------------------------------------------------
struct BexArgs {
    const BexJob* jobs;
    BatchPtr src, dst, other;
    int N;
    const ModConst* mods;
};
__global__ void __launch_bounds__(kElemThreads) k_basis_extend(const __grid_constant__ BexArgs g) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N) return;
    const BexJob& job = g.jobs[blockIdx.y];
    const int b = blockIdx.z;
    const u64* ps = g.src.at(b);
    u64* pd = g.dst.at(b);
    const u64* po = (job.mode == BEX_OTHER_SUB) ? g.other.at(b) : nullptr;
    const size_t N = g.N;
    u64 x[kBexMaxSrc];
#pragma unroll 4
    for (int i = 0; i < job.ns; i++) x[i] = ps[(size_t)job.src_slot[i] * N + j];
    bex_coeff(
        job, g.mods, x, [&](int t) { return po[(size_t)job.other_slot[t] * N + j]; },
        [&](int t, u64 v) { pd[(size_t)job.dst_slot[t] * N + j] = v; });
}
// Fast path: plain CRT extension (ModUp / ModDown of the key-switch) with NS source limbs, every modulus below 2^47.
// Two coefficients per thread (128-bit accesses), the job's constants staged in shared memory as doubles, the CRT on
// the FP64 pipe (lsb_elem_core.h has the same arithmetic in its generic form; tests/emul checks it on the CPU).
template <int NS>
struct BexTargetC {
    double qd, qinvd;
    double vS[NS + 1];
    TwD c[NS];
    int slot, pad_;
};
template <int NS>
__global__ void __launch_bounds__(kElemThreads) k_bex_f64(const __grid_constant__ BexArgs g) {
    __shared__ BexTargetC<NS> sh[kBexMaxDst];
    const BexJob& job = g.jobs[blockIdx.y];
    const int nt = job.nt;
    for (int t = threadIdx.x; t < nt; t += blockDim.x) {
        const ModConst& m = g.mods[job.dst_mod[t]];
        BexTargetC<NS>& c = sh[t];
        c.qd = m.qd;
        c.qinvd = m.qinvd;
        c.slot = job.dst_slot[t];
#pragma unroll
        for (int i = 0; i < NS; i++) c.c[i] = job.qhat_mod_d[t][i];
#pragma unroll
        for (int v = 0; v <= NS; v++) c.vS[v] = f64_from_u64(job.vS[t][v]);
    }
    __syncthreads();
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int b = blockIdx.z;
    const u64* ps = g.src.at(b);
    u64* pd = g.dst.at(b);
    const size_t N = g.N;
    double ya[NS], yb[NS];
    // The overflow count is floor of Lattigo's float64 sum  v = (..(y_0/s_0) + y_1/s_1 ..)  -- divisions, in limb order.
    // A double division costs ~20 FP64-pipe instructions and this kernel is FP64-pipe bound, so the sum is first formed
    // with multiplications by 1/s_i: the two sums differ by less than NS * 2^-51, hence have the same floor unless the
    // fast one lies within 2^-40 of an integer; only then (probability ~2^-39 per coefficient) is the exact sequence run.
    double vma = 0.0, vmb = 0.0;
#pragma unroll
    for (int i = 0; i < NS; i++) {
        const W2 x = ld2(ps + (size_t)job.src_slot[i] * N, j2);
        const ModConst& ms = g.mods[job.src_mod[i]];
        const TwD qi = job.qhat_inv_d[i];
        ya[i] = f64_canon_d(f64_mulmod(f64_from_u64(x.a), qi, ms.qd), ms.qd, ms.qinvd);
        yb[i] = f64_canon_d(f64_mulmod(f64_from_u64(x.b), qi, ms.qd), ms.qd, ms.qinvd);
        vma = d_fma(ya[i], job.rs_f64[i], vma);
        vmb = d_fma(yb[i], job.rs_f64[i], vmb);
    }
    int va = (int)vma, vb = (int)vmb;
    {
        const double eps = 9.094947017729282e-13;  // 2^-40
        const double fa = vma - (double)va, fb = vmb - (double)vb;
        if (fa < eps || fa > 1.0 - eps || fb < eps || fb > 1.0 - eps) {
            double vfa = 0.0, vfb = 0.0;
#pragma unroll
            for (int i = 0; i < NS; i++) {
                const double sf = job.s_f64[i];
                vfa += ya[i] / sf;  // same float64 overflow estimate, same summation order as the generic path
                vfb += yb[i] / sf;
            }
            va = (int)(u64)vfa;
            vb = (int)(u64)vfb;
        }
    }
    for (int t = 0; t < nt; t++) {
        const BexTargetC<NS>& c = sh[t];
        double aa = 0.0, ab = 0.0;
#pragma unroll
        for (int i = 0; i < NS; i++) {
            aa = d_add(aa, f64_mulmod(ya[i], c.c[i], c.qd));
            ab = d_add(ab, f64_mulmod(yb[i], c.c[i], c.qd));
        }
        aa = d_add(aa, -c.vS[va]);
        ab = d_add(ab, -c.vS[vb]);
        W2 o;
        o.a = f64_canon_u(aa, c.qd, c.qinvd);
        o.b = f64_canon_u(ab, c.qd, c.qinvd);
        st2(pd + (size_t)c.slot * N, j2, o);
    }
}

void launch_basis_extend(const DeviceTables& dt, const BexJob* jobs, int njobs, BatchPtr src, BatchPtr dst,
                         BatchPtr other, int batch, cudaStream_t s, const BexMeta* meta) {
    if (batch <= 0 || njobs <= 0) return;
    // consecutive jobs of the same class go in one launch
    for (int j0 = 0; j0 < njobs;) {
        const int fast_ns = (meta && meta[j0].f64_ok && !meta[j0].centered && meta[j0].mode == BEX_PLAIN && meta[j0].ns >= 2 &&
                             meta[j0].ns <= 6)
                                ? meta[j0].ns
                                : 0;
        int j1 = j0 + 1;
        while (j1 < njobs) {
            const int f = (meta && meta[j1].f64_ok && !meta[j1].centered && meta[j1].mode == BEX_PLAIN && meta[j1].ns >= 2 &&
                           meta[j1].ns <= 6)
                              ? meta[j1].ns
                              : 0;
            if (f != fast_ns) break;
            j1++;
        }
        BexArgs g{jobs + j0, src, dst, other, dt.N, dt.mods};
        ProfScope _ps(dt.prof, PK_BASIS_EXT, s, 0.0);
        if (fast_ns) {
            dim3 grid((dt.N / 2 + kElemThreads - 1) / kElemThreads, j1 - j0, batch);
            switch (fast_ns) {
                case 2: k_bex_f64<2><<<grid, kElemThreads, 0, s>>>(g); break;
                case 3: k_bex_f64<3><<<grid, kElemThreads, 0, s>>>(g); break;
                case 4: k_bex_f64<4><<<grid, kElemThreads, 0, s>>>(g); break;
                case 5: k_bex_f64<5><<<grid, kElemThreads, 0, s>>>(g); break;
                default: k_bex_f64<6><<<grid, kElemThreads, 0, s>>>(g); break;
            }
        } else {
            dim3 grid((dt.N + kElemThreads - 1) / kElemThreads, j1 - j0, batch);
            k_basis_extend<<<grid, kElemThreads, 0, s>>>(g);
        }
        LSB_CUDA(cudaGetLastError());
        j0 = j1;
    }
}

// ---- (a - b) * c (+ addend) ------------------------------------------------------------------------
struct FinishArgs {
    BatchPtr a, b, addend, out;
    const FinishMap* fm;
    int N, has_add;
    const ModConst* mods;
};
__global__ void __launch_bounds__(kElemThreads) k_finish(const __grid_constant__ FinishArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int r = blockIdx.y, b = blockIdx.z;
    const FinishMap& fm = *g.fm;
    const u64 q = g.mods[fm.mod[r]].q;
    const Tw c = fm.c[r];
    const W2 x = ld2(g.a.at(b) + (size_t)fm.a_slot[r] * g.N, j2);
    const W2 y = ld2(g.b.at(b) + (size_t)fm.b_slot[r] * g.N, j2);
    W2 o;
    o.a = mul_shoup(submod(x.a, y.a, q), c, q);
    o.b = mul_shoup(submod(x.b, y.b, q), c, q);
    if (g.has_add && fm.add_slot[r] != 0xFF) {
        const W2 z = ld2(g.addend.at(b) + (size_t)fm.add_slot[r] * g.N, j2);
        o.a = addmod(o.a, z.a, q);
        o.b = addmod(o.b, z.b, q);
    }
    st2(g.out.at(b) + (size_t)fm.out_slot[r] * g.N, j2, o);
}
void launch_finish(const DeviceTables& dt, BatchPtr a, BatchPtr b, BatchPtr addend, bool has_addend, BatchPtr out,
                   const FinishMap* fm_dev, int nrows, int batch, cudaStream_t s) {
    if (batch <= 0 || nrows <= 0) return;
    FinishArgs g{a, b, addend, out, fm_dev, dt.N, has_addend ? 1 : 0, dt.mods};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * nrows * batch * 4);
    k_finish<<<elem_grid(dt.N, nrows, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- key-switch inner product -------------------------------------------------------------------------
struct KsInnerK {
    KsInnerArgs a;
    int N;
    const ModConst* mods;
};
template <bool F64, bool MAPPED>
__global__ void __launch_bounds__(kElemThreads) k_ks_inner(const __grid_constant__ KsInnerK g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const KsInnerArgs& a = g.a;
    const int row = blockIdx.y, b = blockIdx.z;
    const int t = MAPPED ? a.tmap[row] : row;
    const int lk = MAPPED ? a.rows : a.l + a.k, klimbs = MAPPED ? a.rows : a.key_level + 1 + a.k;
    const int key_limb = MAPPED ? row : (t < a.l ? t : a.key_level + 1 + (t - a.l));
    const ModConst& mc = g.mods[t < a.l ? t : a.nq + (t - a.l)];
    const size_t N = g.N;
    const u64* pext = a.ext.at(b);
    const u64* pc = a.cntt.at(b) + (size_t)a.cntt_slot0 * N;
    u64* pacc = a.acc.at(b);
    if constexpr (F64) {
        // all primes below 2^47: products and sums on the FP64 pipe (|sum| <= 1.5 * beta * q < 2^51)
        const double q = mc.qd, qinv = mc.qinvd;
        double s0a = 0.0, s0b = 0.0, s1a = 0.0, s1b = 0.0;
        for (int d = 0; d < a.beta; d++) {
            const bool own = t < a.l && t >= d * a.alpha && t < (d + 1) * a.alpha;
            const W2 x = own ? ld2(pc + (size_t)row * N, j2) : ld2(pext + (size_t)(d * lk + row) * N, j2);
            const W2 e0 = ld2(a.evk + ((size_t)(d * 2 + 0) * klimbs + key_limb) * N, j2);
            const W2 e1 = ld2(a.evk + ((size_t)(d * 2 + 1) * klimbs + key_limb) * N, j2);
            const double xa = f64_from_u64(x.a), xb = f64_from_u64(x.b);
            s0a = d_add(s0a, f64_mulmod_var(xa, f64_from_u64(e0.a), q, qinv));
            s0b = d_add(s0b, f64_mulmod_var(xb, f64_from_u64(e0.b), q, qinv));
            s1a = d_add(s1a, f64_mulmod_var(xa, f64_from_u64(e1.a), q, qinv));
            s1b = d_add(s1b, f64_mulmod_var(xb, f64_from_u64(e1.b), q, qinv));
        }
        W2 o0, o1;
        o0.a = f64_canon_u(s0a, q, qinv);
        o0.b = f64_canon_u(s0b, q, qinv);
        o1.a = f64_canon_u(s1a, q, qinv);
        o1.b = f64_canon_u(s1b, q, qinv);
        st2(pacc + (size_t)row * N, j2, o0);
        st2(pacc + (size_t)(lk + row) * N, j2, o1);
        return;
    }
    const BarrettQ bq = mc.bq;
    Acc128 s0a{0, 0}, s0b{0, 0}, s1a{0, 0}, s1b{0, 0};
    int terms = 0;
    for (int d = 0; d < a.beta; d++) {
        const bool own = t < a.l && t >= d * a.alpha && t < (d + 1) * a.alpha;
        const W2 x = own ? ld2(pc + (size_t)row * N, j2) : ld2(pext + (size_t)(d * lk + row) * N, j2);
        const W2 e0 = ld2(a.evk + ((size_t)(d * 2 + 0) * klimbs + key_limb) * N, j2);
        const W2 e1 = ld2(a.evk + ((size_t)(d * 2 + 1) * klimbs + key_limb) * N, j2);
        acc_mad(s0a, x.a, e0.a);
        acc_mad(s0b, x.b, e0.b);
        acc_mad(s1a, x.a, e1.a);
        acc_mad(s1b, x.b, e1.b);
        // reduce_wide needs the sum below 2^(sh+64): fold after `fold` products (61-bit primes with beta > 4 only)
        if (++terms == a.fold && d + 1 < a.beta) {
            s0a = Acc128{reduce_wide(s0a.lo, s0a.hi, bq), 0};
            s0b = Acc128{reduce_wide(s0b.lo, s0b.hi, bq), 0};
            s1a = Acc128{reduce_wide(s1a.lo, s1a.hi, bq), 0};
            s1b = Acc128{reduce_wide(s1b.lo, s1b.hi, bq), 0};
            terms = 1;
        }
    }
    W2 o0, o1;
    o0.a = reduce_wide(s0a.lo, s0a.hi, bq);
    o0.b = reduce_wide(s0b.lo, s0b.hi, bq);
    o1.a = reduce_wide(s1a.lo, s1a.hi, bq);
    o1.b = reduce_wide(s1b.lo, s1b.hi, bq);
    st2(pacc + (size_t)row * N, j2, o0);
    st2(pacc + (size_t)(lk + row) * N, j2, o1);
}
void launch_ks_inner(const DeviceTables& dt, const KsInnerArgs& a, int batch, cudaStream_t s) {
    if (batch <= 0) return;
    KsInnerK g{a, dt.N, dt.mods};
    ProfScope _ps(dt.prof, PK_KS_INNER, s, 8.0 * dt.N * (a.l + a.k) * (batch * (a.beta + 2.0) + 2.0 * a.beta));
    const int rows = a.mapped ? a.rows : a.l + a.k;
    if (rows <= 0) return;
    const dim3 grid = elem_grid(dt.N, rows, batch);
    const bool f64 = a.f64 && a.beta <= 16;
    if (a.mapped) {
        if (f64) k_ks_inner<true, true><<<grid, kElemThreads, 0, s>>>(g);
        else k_ks_inner<false, true><<<grid, kElemThreads, 0, s>>>(g);
    } else {
        if (f64) k_ks_inner<true, false><<<grid, kElemThreads, 0, s>>>(g);
        else k_ks_inner<false, false><<<grid, kElemThreads, 0, s>>>(g);
    }
    LSB_CUDA(cudaGetLastError());
}

// ---- ciphertext (op) integer constant: every limb of every polynomial times c_i, or c_i added to every NTT slot of c0 ----
// (bootstrapping: Chebyshev constants, 2*T^2 - 1, the ScaleUp factor; reference: the constant operations inside
//  CkksBtpContext::bootstrap / regular_bootstrapping_v2, mega_ag_executors_gpu.cu:410-426)
struct ScalarArgs {
    BatchPtr a, out;
    int N, nl, add;  // add: poly 0 gets a + c, the other polynomials are copied
    const ModConst* mods;
    Tw c[64];
};
__global__ void __launch_bounds__(kElemThreads) k_scalar(const __grid_constant__ ScalarArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int r = blockIdx.y, b = blockIdx.z, i = r % g.nl;
    const u64 q = g.mods[i].q;
    W2 x = ld2(g.a.at(b) + (size_t)r * g.N, j2);
    if (g.add) {
        if (r < g.nl) {
            x.a = addmod(x.a, g.c[i].w, q);
            x.b = addmod(x.b, g.c[i].w, q);
        }
    } else {
        x.a = mul_shoup(x.a, g.c[i], q);
        x.b = mul_shoup(x.b, g.c[i], q);
    }
    st2(g.out.at(b) + (size_t)r * g.N, j2, x);
}
void launch_scalar(const DeviceTables& dt, BatchPtr a, BatchPtr out, int npoly, int nl, const Tw* consts, bool add, int batch,
                   cudaStream_t s) {
    if (batch <= 0) return;
    LSB_CHECK(nl <= 64, "scalar: too many limbs");
    ScalarArgs g;
    g.a = a;
    g.out = out;
    g.N = dt.N;
    g.nl = nl;
    g.add = add ? 1 : 0;
    g.mods = dt.mods;
    for (int i = 0; i < nl; i++) g.c[i] = consts[i];
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * npoly * nl * batch * 2);
    k_scalar<<<elem_grid(dt.N, npoly * nl, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- rescale ---------------------------------------------------------------------------------------------
struct RescaleLiftArgs {
    BatchPtr last, tmp;
    int N, npoly, level, last_ps, raise;
    const ModConst* mods;
};
// last: coefficient-domain limb of each polynomial (block [npoly] with stride last_ps limbs), residues of the SOURCE
// modulus: q_level for the rescale (targets q_0..q_{level-1}), q_0 for ModRaise (raise = 1: targets q_1..q_level, the
// source limb is kept as limb 0 of the result).  Each target gets the CENTRED lift reduced: ((x + half) mod q_src
// reduced mod q_i) - half mod q_i.
template <bool F64>
__global__ void __launch_bounds__(kElemThreads) k_rescale_lift(const __grid_constant__ RescaleLiftArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int p = blockIdx.y, b = blockIdx.z;
    const u64 ql = g.mods[g.raise ? 0 : g.level].q, half = (ql - 1) >> 1;
    const W2 x0 = ld2(g.last.at(b) + (size_t)p * g.last_ps * g.N, j2);
    W2 x;
    x.a = addmod(x0.a, half, ql);
    x.b = addmod(x0.b, half, ql);
    const int i0 = g.raise ? 1 : 0, i1 = g.raise ? g.level + 1 : g.level;
    u64* pt = g.tmp.at(b) + (size_t)p * i1 * g.N;
    if (g.raise) st2(pt, j2, x0);
    if constexpr (F64) {  // every prime below 2^47: the per-target reductions on the FP64 pipe (integer version: IMAD-bound)
        const double xa = f64_from_u64(x.a), xb = f64_from_u64(x.b);
        for (int i = i0; i < i1; i++) {
            const ModConst& m = g.mods[i];
            const double hq = f64_from_u64(barrett64(half, m.bq));  // CTA-uniform
            W2 o;
            o.a = f64_canon_u(d_add(f64_centre(xa, m.qd, m.qinvd), -hq), m.qd, m.qinvd);
            o.b = f64_canon_u(d_add(f64_centre(xb, m.qd, m.qinvd), -hq), m.qd, m.qinvd);
            st2(pt + (size_t)i * g.N, j2, o);
        }
        return;
    }
    for (int i = i0; i < i1; i++) {
        const ModConst& m = g.mods[i];
        const u64 hq = barrett64(half, m.bq);  // CTA-uniform
        W2 o;
        o.a = submod(barrett64(x.a, m.bq), hq, m.q);
        o.b = submod(barrett64(x.b, m.bq), hq, m.q);
        st2(pt + (size_t)i * g.N, j2, o);
    }
}
void launch_rescale_lift(const DeviceTables& dt, BatchPtr last, int last_poly_stride, BatchPtr tmp, int npoly, int level,
                         int batch, cudaStream_t s, bool raise) {
    if (batch <= 0) return;
    RescaleLiftArgs g{last, tmp, dt.N, npoly, level, last_poly_stride, raise ? 1 : 0, dt.mods};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * npoly * batch * (1 + level));
    bool f64 = elem_f64_enabled();
    for (int i = 0; i <= level; i++) f64 = f64 && dt.elem_f64_ok[i];
    if (f64) k_rescale_lift<true><<<elem_grid(dt.N, npoly, batch), kElemThreads, 0, s>>>(g);
    else k_rescale_lift<false><<<elem_grid(dt.N, npoly, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

struct BfvRescaleArgs {
    BatchPtr in, out;
    int N, npoly, level;
    const Tw* qlinv;  // [level] (ql^-1 mod q_i)
    const ModConst* mods;
};
__global__ void __launch_bounds__(kElemThreads) k_bfv_rescale(const __grid_constant__ BfvRescaleArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int p = blockIdx.y, b = blockIdx.z;
    const int l = g.level + 1;
    const u64 ql = g.mods[g.level].q, half = (ql - 1) >> 1;
    const u64* pin = g.in.at(b) + (size_t)p * l * g.N;
    u64* pout = g.out.at(b) + (size_t)p * (l - 1) * g.N;
    W2 x = ld2(pin + (size_t)g.level * g.N, j2);
    x.a = addmod(x.a, half, ql);
    x.b = addmod(x.b, half, ql);
    for (int i = 0; i < g.level; i++) {
        const ModConst& m = g.mods[i];
        const u64 hq = barrett64(half, m.bq);
        const Tw c = g.qlinv[i];
        const W2 sv = ld2(pin + (size_t)i * g.N, j2);
        W2 o;
        o.a = mul_shoup(submod(sv.a, submod(barrett64(x.a, m.bq), hq, m.q), m.q), c, m.q);
        o.b = mul_shoup(submod(sv.b, submod(barrett64(x.b, m.bq), hq, m.q), m.q), c, m.q);
        st2(pout + (size_t)i * g.N, j2, o);
    }
}
void launch_bfv_rescale(const DeviceTables& dt, BatchPtr in, BatchPtr out, int npoly, int level, const Tw* qlinv_dev,
                        int batch, cudaStream_t s) {
    if (batch <= 0) return;
    BfvRescaleArgs g{in, out, dt.N, npoly, level, qlinv_dev, dt.mods};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * npoly * batch * (2 * level + 1));
    k_bfv_rescale<<<elem_grid(dt.N, npoly, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- automorphisms -----------------------------------------------------------------------------------------
struct PermArgs {
    BatchPtr in, out;
    int N, logN;
    const u32* index;
    u64 gal;
    const ModConst* mods;
    RowMap map;
};
__global__ void __launch_bounds__(kElemThreads) k_permute_ntt(const __grid_constant__ PermArgs g) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N) return;
    const int r = blockIdx.y, b = blockIdx.z;
    const u64* pi = g.in.at(b) + (size_t)g.map.in_slot[r] * g.N;
    u64* po = g.out.at(b) + (size_t)g.map.out_slot[r] * g.N;
    po[j] = pi[g.index[j]];
}
void launch_permute_ntt(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, const u32* index,
                        int batch, cudaStream_t s) {
    if (batch <= 0 || map.n <= 0) return;
    PermArgs g{in, out, dt.N, dt.logN, index, 0, dt.mods, map};
    dim3 grid((dt.N + kElemThreads - 1) / kElemThreads, map.n, batch);
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * map.n * batch * 2);
    k_permute_ntt<<<grid, kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}
// ring.Permute: X^i -> X^(i*gal); scatter form out[(i*gal) mod N] = +-in[i]
__global__ void __launch_bounds__(kElemThreads) k_permute_coeff(const __grid_constant__ PermArgs g) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N) return;
    const int r = blockIdx.y, b = blockIdx.z;
    const u64 q = g.mods[g.map.mod[r]].q;
    const u64* pi = g.in.at(b) + (size_t)g.map.in_slot[r] * g.N;
    u64* po = g.out.at(b) + (size_t)g.map.out_slot[r] * g.N;
    const u64 idx = (u64)j * g.gal;
    const u64 v = pi[j];
    po[idx & (u64)(g.N - 1)] = ((idx >> g.logN) & 1) ? negmod(v, q) : v;
}
void launch_permute_coeff(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, u64 gal, int batch,
                          cudaStream_t s) {
    if (batch <= 0 || map.n <= 0) return;
    PermArgs g{in, out, dt.N, dt.logN, nullptr, gal, dt.mods, map};
    dim3 grid((dt.N + kElemThreads - 1) / kElemThreads, map.n, batch);
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * map.n * batch * 2);
    k_permute_coeff<<<grid, kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

// ---- plaintext helpers -----------------------------------------------------------------------------------------
struct LiftArgs {
    BatchPtr in, out;
    int N, nl;
};
__global__ void __launch_bounds__(kElemThreads) k_lift_ringt(const __grid_constant__ LiftArgs g) {
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= g.N / 2) return;
    const int b = blockIdx.z;
    const W2 x = ld2(g.in.at(b), j2);
    u64* po = g.out.at(b);
    for (int i = 0; i < g.nl; i++) st2(po + (size_t)i * g.N, j2, x);
}
void launch_lift_ringt(const DeviceTables& dt, BatchPtr in, BatchPtr out, int nl, int batch, cudaStream_t s) {
    if (batch <= 0) return;
    LiftArgs g{in, out, dt.N, nl};
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * batch * (1 + nl));
    k_lift_ringt<<<elem_grid(dt.N, 1, batch), kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

struct ScaleUpArgs {
    BatchPtr in, out;
    int N, nl;
    const ScaleUpConsts* sc;
    const ModConst* mods;
};
__global__ void __launch_bounds__(kElemThreads) k_bfv_scale_up(const __grid_constant__ ScaleUpArgs g) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N) return;
    const int b = blockIdx.z;
    const ScaleUpConsts& sc = *g.sc;
    const u64 m = barrett64(g.in.at(b)[j], sc.bt);
    const u64 r = addmod(mulmod(m, sc.qmodt, sc.bt), sc.thalf, sc.t);
    u64* po = g.out.at(b);
    for (int i = 0; i < g.nl; i++) {
        const ModConst& mc = g.mods[i];
        const u64 v = submod(barrett64(r, mc.bq), sc.thalf_q[i], mc.q);
        po[(size_t)i * g.N + j] = mul_shoup(v, sc.neg_tinv[i], mc.q);
    }
}
void launch_bfv_scale_up(const DeviceTables& dt, BatchPtr in, BatchPtr out, int nl, const ScaleUpConsts* sc_dev,
                         int batch, cudaStream_t s) {
    if (batch <= 0) return;
    ScaleUpArgs g{in, out, dt.N, nl, sc_dev, dt.mods};
    dim3 grid((dt.N + kElemThreads - 1) / kElemThreads, 1, batch);
    ProfScope _ps(dt.prof, PK_ELEMWISE, s, 8.0 * dt.N * batch * (1 + nl));
    k_bfv_scale_up<<<grid, kElemThreads, 0, s>>>(g);
    LSB_CUDA(cudaGetLastError());
}

}  // namespace lsb
