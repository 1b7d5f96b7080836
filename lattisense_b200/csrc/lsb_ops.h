// Ciphertext-level operations of the mega_ag executor, batched: every call processes `batch`
// independent ciphertexts with the same shape (the runner groups identical DAG nodes; the
// benchmark drives 1024-ciphertext batches).  Layouts follow the reference ABI
// (abi/c_types.h:26-60 flattened): ct = [degree+1][level+1][N] u64, evk = [beta][2][key_level+1+k][N],
// CKKS data in the NTT domain, BFV data in the coefficient domain.
#pragma once
#include "lsb_context.h"

namespace lsb {

struct OpEnv {
    Context& c;
    Workspace& ws;
    cudaStream_t s;
    int chunk;  // max batch items per pass through the scratch buffers (0 = auto)
};

// bind_gpu_mult (mega_ag_executors_gpu.cu:177-227): ct x ct -> degree-2 ciphertext
void op_mult(OpEnv& e, int level, BatchPtr a, BatchPtr b, BatchPtr out3, int batch);
// bind_gpu_relin (:229-238)
void op_relin(OpEnv& e, int level, BatchPtr ct3, const u64* rlk, int key_level, BatchPtr out2, int batch);
// bind_gpu_rescale (:240-248); CKKS divide-and-round by q_level, BFV modulus switch
void op_rescale(OpEnv& e, int level, int npoly, BatchPtr in, BatchPtr out, int batch);
// bind_gpu_drop_level (:250-262)
void op_mod_raise(OpEnv& e, int top_level, int npoly, BatchPtr in, BatchPtr out, int batch);
void op_linear_transform(OpEnv& e, int level, const u64* ct, int n_diag, const int* diag_rot, const u64* const* diag_pt, int n1,
                         int n_keys, const int* key_rot, const u64* const* glk, int key_level, u64* out);
void op_drop_level(OpEnv& e, int level, int npoly, BatchPtr in, BatchPtr out, int batch);
void op_drop_to(OpEnv& e, int level_from, int level_to, int npoly, BatchPtr in, BatchPtr out, int batch);
// key switch without automorphism (bootstrapping's swk_dts / swk_std) and ct (op) integer constant
void op_switch_key(OpEnv& e, int level, BatchPtr ct, const u64* swk, int key_level, BatchPtr out, int batch);
void op_scalar(OpEnv& e, int level, int npoly, BatchPtr ct, const u64* residues, bool add, BatchPtr out, int batch);
// bind_gpu_rotate_col / bind_gpu_rotate_row (:264-292): automorphism gal + key-switch under glk[gal]
void op_rotate(OpEnv& e, int level, BatchPtr ct, const u64* glk, int key_level, u64 gal, BatchPtr out, int batch);
// several rotations of the same ciphertexts sharing one decomposition (hoisting); limbs equal op_rotate's
void op_rotate_hoisted(OpEnv& e, int level, BatchPtr ct, int n_rot, const u64* const* glk, int key_level, const u64* gal, const BatchPtr* out,
                       int batch);
// bind_gpu_add / sub / neg (:71-175), ct (op) ct with npoly polynomials
void op_add(OpEnv& e, int level, int npoly, BatchPtr a, BatchPtr b, BatchPtr out, int batch);
void op_sub(OpEnv& e, int level, int npoly, BatchPtr a, BatchPtr b, BatchPtr out, int batch);
void op_neg(OpEnv& e, int level, int npoly, BatchPtr a, BatchPtr out, int batch);
// ct (op) pt : pt = [level+1][N] in the scheme's native domain (added to / subtracted from c0)
void op_add_plain(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, bool subtract, int batch);
// ct (op) pt_ringt : pt = [1][N] coefficients mod t (BFV: ScaleUp then add; CKKS: not defined by the reference CPU path)
void op_add_plain_ringt(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, bool subtract,
                        int batch);
// ct x pt_mul (NTT-domain plaintext, Montgomery already removed: mf_nbits = 0)
void op_mul_plain(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, int batch);
// ct x pt_ringt
void op_mul_plain_ringt(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, int batch);

// bind_gpu_cmp_sum / bind_gpu_cmpac_sum (:294-408): sum_i ct_i x pt_i (+ acc_in); pts are pt_ringt ([1][N] mod t)
// when `ringt`, NTT-domain [level+1][N] plaintexts otherwise
void op_sum_terms(OpEnv& e, int level, int npoly, const u64* const* ct_tab, const u64* const* pt_tab, u64* const* out_tab,
                  int n_mac, int n_plain, int batch);
void op_mac(OpEnv& e, int level, int nterms, const BatchPtr* cts, const BatchPtr* pts, bool ringt, bool has_acc,
            BatchPtr acc_in, BatchPtr out, int batch);

// fused paths used by the runner's pattern matcher and by the benchmark
void op_mult_relin(OpEnv& e, int level, BatchPtr a, BatchPtr b, const u64* rlk, int key_level, BatchPtr out2,
                   int batch);
void op_mult_relin_rescale(OpEnv& e, int level, BatchPtr a, BatchPtr b, const u64* rlk, int key_level, BatchPtr out2,
                           int batch);

// limb-sharded key-switch phases (lsb_shard.cu; the all-gathers between them are the caller's)
struct ShardPlan;
// `peers` (optional, world pointers): fused all-gather -- the phase's INTT stores into every rank's gathered buffer
void op_ks_shard_begin(OpEnv& e, const ShardPlan& sp, BatchPtr c_own_ntt, BatchPtr coeff_own, int batch,
                       u64* const* peers = nullptr);
void op_ks_shard_inner(OpEnv& e, const ShardPlan& sp, BatchPtr coeff_all, BatchPtr c_own_ntt, const u64* evk_own,
                       BatchPtr acc_own, BatchPtr accP_own, int batch, u64* const* peers = nullptr);
void op_ks_shard_end(OpEnv& e, const ShardPlan& sp, BatchPtr accP_all, BatchPtr acc_own, BatchPtr addend_own, bool has_addend,
                     BatchPtr out_own, int batch);

// plain batched NTT over [rows][N] limbs whose moduli are mod_idx[r % period]
void op_ntt(OpEnv& e, BatchPtr in, BatchPtr out, int rows, const int* mod_idx, int period, bool inverse, int batch);

}  // namespace lsb
