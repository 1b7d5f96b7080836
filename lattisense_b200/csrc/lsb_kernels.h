// Kernel launchers of the B200 executor core (all asynchronous on the given stream).
#pragma once
#include "lsb_common.h"
#include "lsb_prof.h"

namespace lsb {

// Device-resident tables of one context (one parameter set on one GPU).
struct DeviceTables {
    int device = 0;
    int sm_count = 148;
    int logN = 0, N = 0;
    int nmod = 0;
    const Tw* tw_fwd = nullptr;  // [nmod][N]  psi^bitrev(i) with Shoup companion
    const Tw* tw_inv = nullptr;  // [nmod][N]  psi^-bitrev(i)
    const TwD* tw_fwd_d = nullptr;  // FP64 policy: (w, w/q) pairs, zero for ineligible primes
    const TwD* tw_inv_d = nullptr;
    std::vector<uint8_t> elem_f64_ok;  // per modulus: below 2^47, element-wise kernels may use the FP64 pipe
    // the same tables with the stages of the cluster kernel's last round in tail-SoA order (lsb_ntt_arith.h)
    const Tw* twc_fwd = nullptr;
    const Tw* twc_inv = nullptr;
    const TwD* twc_fwd_d = nullptr;
    const TwD* twc_inv_d = nullptr;
    const double* twc_fwd_w = nullptr;  // the same tables, w only (ArithF64W)
    const double* twc_inv_w = nullptr;
    const ModConst* mods = nullptr;
    Profiler* prof = nullptr;
    std::vector<uint8_t> f64_ok;  // host copy: modulus qualifies for the FP64 butterfly policy
    // side stream + events on which the FP64-policy rows of a mixed transform run beside the integer-policy rows
    // (launch_ntt_cluster); owned by the Context
    cudaStream_t ntt_side = nullptr;
    cudaEvent_t ntt_fork = nullptr, ntt_join = nullptr;
};

// ---- NTT ----
void launch_ntt(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch, bool inverse,
                cudaStream_t s);

// cluster kernel (lsb_ntt_cluster.cu); launch_ntt dispatches to it unless LSB_NTT_IMPL=single
struct NttFinish;
void launch_ntt_cluster(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, int batch, bool inverse,
                        const std::vector<uint8_t>& f64_ok, bool allow_f64, cudaStream_t s, const NttFinish* fin = nullptr,
                        u64* const* peers = nullptr, int npeer = 0);
// inverse NTT whose outputs go to word offset map.out_slot[r] * N of each of the `npeer` buffers (peer-mapped memory of
// other GPUs included): the limb-sharded key-switch's all-gather fused into the transform that feeds it
void launch_intt_to_peers(const DeviceTables& dt, BatchPtr in, const RowMap& map, u64* const* peers, int npeer, cudaStream_t s);

// ---- basis extension (ModUp / ModDown / BFV scaling), Lattigo-style with float64 overflow count ----
constexpr int kBexMaxSrc = 16;
constexpr int kBexMaxDst = 48;
enum BexMode {
    BEX_PLAIN = 0,      // dst = ext
    BEX_OTHER_SUB = 1,  // dst = (other - ext) * mulc + addc
    BEX_SUB_MUL = 2,    // dst = (ext - addc) * mulc
};
struct BexJob {
    int ns, nt, mode;
    int centered;  // ns == 1 in key-switch ModUp: centred lift instead of the CRT formula
    int wide_ok;   // ns * max(source modulus) < 2^63: the CRT sums fit reduce_wide (else two-word Barrett)
    int f64_ok;    // every modulus involved is below 2^47: the CRT runs on the FP64 pipe (tables below)
    uint8_t src_slot[kBexMaxSrc], src_mod[kBexMaxSrc];
    uint8_t dst_slot[kBexMaxDst], dst_mod[kBexMaxDst], other_slot[kBexMaxDst];
    Tw qhat_inv[kBexMaxSrc];
    double s_f64[kBexMaxSrc];
    double rs_f64[kBexMaxSrc];  // 1 / s_f64 (fast overflow estimate; the exact division is the fallback near integers)
    u64 qhat_mod[kBexMaxDst][kBexMaxSrc];
    u64 vS[kBexMaxDst][kBexMaxSrc + 1];
    Tw mulc[kBexMaxDst];
    u64 addc[kBexMaxDst];
    // FP64 copies (valid when f64_ok): (w, w/q) pairs with respect to the modulus the product is reduced by
    TwD qhat_inv_d[kBexMaxSrc];
    TwD qhat_mod_d[kBexMaxDst][kBexMaxSrc];
    TwD mulc_d[kBexMaxDst];
    // Mixed-pipe CRT (integer path, !f64_ok): targets flagged in t_f64 (modulus below 2^47) are summed on the FP64 pipe while
    // the others use the 64-bit integer multiplier, so that both multiply pipes of an SM work on one job.  qhat_mod_d[t] /
    // mulc_d[t] are then valid for the flagged targets; when a source modulus reaches 2^51 (src_split) y_i enters as
    // y_hi * 2^32 + y_lo and qhat_mod_hi_d[t][i] holds (qhat_mod[t][i] * 2^32 mod p_t).
    int n_f64, src_split;
    uint8_t t_f64[kBexMaxDst];
    TwD qhat_mod_hi_d[kBexMaxDst][kBexMaxSrc];
};
// jobs: device array of njobs BexJob; src/dst/other: per-batch-item blocks.
// meta: host-side description of the jobs (selects the specialised kernels); nullptr: generic kernel
struct BexMeta {
    int ns, nt, mode, centered, f64_ok;
};
void launch_basis_extend(const DeviceTables& dt, const BexJob* jobs, int njobs, BatchPtr src, BatchPtr dst,
                         BatchPtr other, int batch, cudaStream_t s, const BexMeta* meta = nullptr);

// ---- limb-wise kernels ----
enum BinOp { OP_ADD = 0, OP_SUB = 1 };
// out[r] = a[r] (op) b[r] for r < map.n ; rows r >= b_rows copy a (or negate-free pass through)
void launch_binary(const DeviceTables& dt, BinOp op, BatchPtr a, BatchPtr b, BatchPtr out, const RowMap& map,
                   int b_rows, int batch, cudaStream_t s);
void launch_negate(const DeviceTables& dt, BatchPtr a, BatchPtr out, const RowMap& map, int batch, cudaStream_t s);
// degree-1 x degree-1 tensor over nl limbs: a,b blocks [2][a_stride_limbs] -> out [3][nl]
void launch_tensor(const DeviceTables& dt, BatchPtr a, BatchPtr b, BatchPtr out, int nl, int a_poly_stride,
                   int out_poly_stride, const uint8_t* mod_of_limb, int batch, cudaStream_t s);
// out[p][i] (+)= ct[p][i] * pt[i]  (NTT domain), p < npoly
// how many term ranges a sum should be cut into so that the launch fills the machine (1: no split)
int sum_terms_split(const DeviceTables& dt, int npoly, int nl, int n_terms, int batch);
void launch_sum_terms(const DeviceTables& dt, const u64* const* ct, const u64* const* pt, u64* const* out, int npoly, int nl,
                      int n_mac, int n_plain, int batch, cudaStream_t s, int split = 1, u64* partial = nullptr);
void launch_mul_plain(const DeviceTables& dt, BatchPtr ct, BatchPtr pt, BatchPtr out, int npoly, int nl,
                      bool accumulate, int batch, cudaStream_t s);
// out[r] = (a[a_slot[r]] - b[b_slot[r]]) * c_r (+ addend[add_slot[r]])   (ModDown / rescale finish)
struct FinishMap {
    int n;
    uint8_t a_slot[kMaxRows], b_slot[kMaxRows], add_slot[kMaxRows], out_slot[kMaxRows], mod[kMaxRows];
    Tw c[kMaxRows];  // multiplier of row r (Shoup pair w.r.t. the row's modulus)
};
void launch_finish(const DeviceTables& dt, BatchPtr a, BatchPtr b, BatchPtr addend, bool has_addend, BatchPtr out,
                   const FinishMap* fm_dev, int nrows, int batch, cudaStream_t s);
// finish fused into the forward NTT that produces its `b` operand (cluster kernel epilogue): the transform of
// b[fm.b_slot[r]] never goes to memory.  fm_host is the host copy of *fm_dev.
struct NttFinish {
    const FinishMap* fm_dev;
    const FinishMap* fm_host;
    BatchPtr a, addend, out;
    bool has_addend;
};
void launch_ntt_finish(const DeviceTables& dt, BatchPtr b, const NttFinish& fin, int batch, cudaStream_t s);

// key-switch inner product: acc[p][t] = sum_d D_d[t] * evk[d][p][keylimb(t)]
struct KsInnerArgs {
    BatchPtr ext;    // [beta][l+k] limbs (own-digit slots unused)
    BatchPtr cntt;   // NTT-domain input polynomial
    int cntt_slot0;  // limb t of the input polynomial is at slot cntt_slot0 + t
    const u64* evk;  // [beta_key][2][key_level+1+k][N]
    BatchPtr acc;    // [2][l+k]
    int l, k, beta, alpha, key_level;
    int nq;  // modulus index of P_0
    int f64;   // every prime involved is below 2^47 (FP64-pipe variant)
    // limb-sharded variant (lsb_shard.cu): only `rows` of the l+k limbs are present, compacted; row r is global limb
    // tmap[r] (Q index, or l + P index); ext / acc / evk / cntt are all indexed by the compact row
    int mapped, rows;
    uint8_t tmap[64];
    int fold;  // products that may be accumulated before a reduction: fold * max(q) < 2^63 (reduce_wide bound)
};
void launch_ks_inner(const DeviceTables& dt, const KsInnerArgs& a, int batch, cudaStream_t s);

// ct (op) per-limb integer constant (Shoup pairs): multiply every polynomial, or add to every slot of polynomial 0
void launch_scalar(const DeviceTables& dt, BatchPtr a, BatchPtr out, int npoly, int nl, const Tw* consts, bool add, int batch,
                   cudaStream_t s);

// CKKS rescale lift: tmp[p][i] = ((last[p] + half) mod q_i) - half mod q_i, i < level
void launch_rescale_lift(const DeviceTables& dt, BatchPtr last, int last_poly_stride, BatchPtr tmp, int npoly, int level,
                         int batch, cudaStream_t s, bool raise = false);
// BFV rescale (coefficient domain, fully fused): [np][l] -> [np][l-1]
void launch_bfv_rescale(const DeviceTables& dt, BatchPtr in, BatchPtr out, int npoly, int level, const Tw* qlinv_dev,
                        int batch, cudaStream_t s);

// automorphisms over rows of map: NTT domain gather (index table) / coefficient domain signed permutation
void launch_permute_ntt(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, const u32* index,
                        int batch, cudaStream_t s);
void launch_permute_coeff(const DeviceTables& dt, BatchPtr in, BatchPtr out, const RowMap& map, u64 gal, int batch,
                          cudaStream_t s);
// lift a single-limb plaintext mod t (pt_ringt) onto nl limbs (values in [0,t) are copied as they are)
void launch_lift_ringt(const DeviceTables& dt, BatchPtr in, BatchPtr out, int nl, int batch, cudaStream_t s);
// BFV ScaleUp of a mod-t plaintext: out_i = round(Q*m/t) mod q_i
struct ScaleUpConsts {
    u64 t, thalf, qmodt;
    BarrettQ bt;
    u64 thalf_q[64];
    Tw neg_tinv[64];
};
void launch_bfv_scale_up(const DeviceTables& dt, BatchPtr in, BatchPtr out, int nl, const ScaleUpConsts* sc_dev,
                         int batch, cudaStream_t s);

}  // namespace lsb
