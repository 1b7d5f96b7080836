/*
 * lsb200.h — kernel-level C ABI of the B200 executor core (libgpu_mega_ag_runner.so).
 *
 * This is the thin C layer the host runtime (and the reference-facing runner in
 * lattisense_gpu_runner.h) calls CUDA through.  Each entry point names the reference interface it
 * stands in for; paths are relative to the reference tree (cipherflow-fhe/lattisense @ 88bf833).
 *
 * Conventions (reference ABI, abi/c_types.h:26-60 flattened into contiguous device arrays):
 *   - residues are uint64_t in [0, q_i); limbs ordered q_0..q_level, keys then p_0..p_{k-1}
 *     (plug-in/lattigo/acc/c_struct_import_export.go:41-59)
 *   - ciphertext  : [degree+1][level+1][N]
 *   - eval key    : [beta][2][key_level+1+k][N], beta = ceil((key_level+1)/k), NTT domain, Montgomery removed
 *   - CKKS ciphertexts/plaintexts are in the (Lattigo) NTT domain, BFV ones in the coefficient domain
 *   - a batch is `batch` such blocks at a constant stride (`*_stride` words; 0 = densely packed)
 *   - all pointers named d_* are DEVICE pointers on the context's GPU; `stream` is a cudaStream_t
 *     (NULL = default stream).  Calls are asynchronous on that stream unless stated otherwise.
 *   - every function returning int returns 0 on success, non-zero on failure (message via
 *     lsb_last_error); there is no CPU fallback: without a CUDA device context creation fails.
 */
#ifndef LSB200_H
#define LSB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lsb_context_st* lsb_context_t;

enum { LSB_SCHEME_BFV = 0, LSB_SCHEME_CKKS = 1 };

/* Last error message of the calling thread (empty string if none). */
const char* lsb_last_error(void);
/* Library version string. */
const char* lsb_version(void);

/* init_gpu_context (mega_ag_runners/gpu/gpu_wrapper.cu:53-138): modulus chain Q = q[0..nq), P = p[0..np),
 * plaintext modulus t (BFV; 0 for CKKS), NTT tables.  Built once and reused (the reference rebuilds per run). */
lsb_context_t lsb_context_create(int device, int scheme, int log_n, const uint64_t* q, int nq, const uint64_t* p,
                                 int np, uint64_t t);
void lsb_context_destroy(lsb_context_t ctx);
int lsb_context_num_moduli(lsb_context_t ctx);           /* Q, P and (BFV) the auxiliary QMul basis */
uint64_t lsb_context_modulus(lsb_context_t ctx, int i);
int lsb_context_level_qmul(lsb_context_t ctx, int level); /* BFV: index of the last QMul prime used at `level` */

/* Device memory + transfers (gpu_abi_bridge_executors.h:60-191 does one cudaMemcpyAsync per limb). */
uint64_t* lsb_device_alloc(lsb_context_t ctx, size_t words);
void lsb_device_free(lsb_context_t ctx, uint64_t* d_ptr);
int lsb_upload(lsb_context_t ctx, uint64_t* d_dst, const uint64_t* h_src, size_t words, void* stream);
int lsb_download(lsb_context_t ctx, uint64_t* h_dst, const uint64_t* d_src, size_t words, void* stream);
int lsb_stream_sync(lsb_context_t ctx, void* stream);
/* Max batch items per scratch pass (0 = automatic, sized to keep scratch L2-resident). */
void lsb_set_chunk(lsb_context_t ctx, int chunk);

/* Streams: every operation below is enqueued on `stream` and returns without waiting.  A context owns ONE scratch block
 * shared by all its operations; calls are serialised on the host, and when a call names another stream than the previous
 * call on the same context it is first ordered (event wait) after the previous stream's work.  Concurrent GPU execution
 * of two operations therefore needs two contexts. */
/* Batched negacyclic NTT / INTT over `rows` limbs per item; limb r uses modulus index mod_idx[r % period].
 * Stands in for the HEonGPU NTT under every operator; ordering = Lattigo ring.NTT / ring.InvNTT. */
int lsb_ntt(lsb_context_t ctx, const uint64_t* d_in, uint64_t* d_out, int rows, const int* mod_idx, int period,
            int inverse, int batch, void* stream);

/* bind_gpu_mult, ct x ct (mega_ag_executors_gpu.cu:177-227; CPU twin mega_ag_executors_cpu.cpp:186,230). */
int lsb_mult(lsb_context_t ctx, int level, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out3, int batch,
             void* stream);
/* bind_gpu_relin (mega_ag_executors_gpu.cu:229-238; CPU :236-242). */
int lsb_relin(lsb_context_t ctx, int level, const uint64_t* d_ct3, const uint64_t* d_rlk, int key_level,
              uint64_t* d_out2, int batch, void* stream);
/* bind_gpu_rescale (mega_ag_executors_gpu.cu:240-248; CPU :244-255). */
int lsb_rescale(lsb_context_t ctx, int level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                void* stream);
/* CKKS ModRaise, the first step of bootstrapping (reference: CkksBtpContext::bootstrap behind
 * mega_ag_executors_gpu.cu:410-426 / mega_ag_executors_cpu.cpp:451-468; Lattigo ckks/bootstrapping): level-0
 * ciphertext [npoly][1][N] -> [npoly][top_level+1][N], the centred residues mod q_0 re-read modulo every prime. */
int lsb_mod_raise(lsb_context_t ctx, int top_level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                  void* stream);
/* lsb_mod_raise without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same kernels and bit-identical output as lsb_mod_raise; the INTT's intermediate
 * (npoly * N * batch words) lives in a file-scope `static __device__` array of LSB_STATIC_SCRATCH_WORDS 64-bit words
 * (4 MiB, held by every GPU that loads the library until the process exits) instead of the context's workspace.
 * Returns 1 with a message in lsb_last_error() when npoly * N * batch > LSB_STATIC_SCRATCH_WORDS (split the batch, or
 * call lsb_mod_raise).  The call itself makes no cudaMalloc / cudaFree / stream-ordered allocation; its first use on a
 * device resolves the array's address with cudaGetSymbolAddress.  All calls on a device share the array, which keeps the
 * previous call's contents (ModRaise overwrites what it reads, nothing is cleared): calls whose kernels overlap on the
 * GPU -- different streams with no ordering between them -- race; keep them ordered. */
#define LSB_STATIC_SCRATCH_WORDS (1u << 19)
int lsb_mod_raise_static(lsb_context_t ctx, int top_level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                         void* stream);
/* Diagonal-form linear transform, baby-step / giant-step: the CoeffsToSlots / SlotsToCoeffs matrices of CKKS
 * bootstrapping (rotation sets: frontend/bootstrap_params.py:159-209).  diag_rot[i] = diagonal index k_i;
 * d_diag_pt[i] = its plaintext [level+1][N], NTT domain, slot values rotated by -(k_i - k_i mod n1) at encode time;
 * key_rot / d_glk = the column-rotation keys at hand (rotation amount -> key, all at key_level).
 * out = rescale(sum_g rot(sum_b pt_{g+b} * rot(ct, b), g)) at level-1.  One ciphertext per call. */
int lsb_linear_transform(lsb_context_t ctx, int level, const uint64_t* d_ct, int n_diag, const int* diag_rot,
                         const uint64_t* const* d_diag_pt, int n1, int n_keys, const int* key_rot,
                         const uint64_t* const* d_glk, int key_level, uint64_t* d_out, void* stream);
/* bind_gpu_drop_level (mega_ag_executors_gpu.cu:250-262). */
int lsb_drop_level(lsb_context_t ctx, int level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                   void* stream);
/* bind_gpu_rotate_col / bind_gpu_rotate_row (mega_ag_executors_gpu.cu:264-292): automorphism X -> X^gal
 * (gal = 5^step mod 2N for columns, 2N-1 for rows / conjugate; frontend/custom_task.py:139-145) + key-switch. */
int lsb_rotate(lsb_context_t ctx, int level, const uint64_t* d_ct, const uint64_t* d_glk, int key_level,
               uint64_t galois_element, uint64_t* d_out, int batch, void* stream);
/* n_rot rotations of the same ciphertexts with ONE shared decomposition (hoisted rotations, the baby steps of a BSGS linear
 * transform: frontend/custom_task.py:1427-1457): d_out[r] = lsb_rotate(d_ct, d_glk[r], galois_elements[r]) bit for bit. */
int lsb_rotate_hoisted(lsb_context_t ctx, int level, const uint64_t* d_ct, int n_rot, const uint64_t* const* d_glk, int key_level,
                       const uint64_t* galois_elements, uint64_t* const* d_out, int batch, void* stream);
/* bind_gpu_add / sub / neg (mega_ag_executors_gpu.cu:71-175). */
int lsb_add(lsb_context_t ctx, int level, int npoly, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out,
            int batch, void* stream);
int lsb_sub(lsb_context_t ctx, int level, int npoly, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out,
            int batch, void* stream);
int lsb_neg(lsb_context_t ctx, int level, int npoly, const uint64_t* d_a, uint64_t* d_out, int batch, void* stream);
/* ct +/- pt, ct +/- pt_ringt, ct x pt_mul, ct x pt_ringt (plain branches, mega_ag_executors_gpu.cu:92-161,189-214). */
int lsb_add_plain(lsb_context_t ctx, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                  uint64_t* d_out, int subtract, int batch, void* stream);
int lsb_add_plain_ringt(lsb_context_t ctx, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                        uint64_t* d_out, int subtract, int batch, void* stream);
int lsb_mul_plain(lsb_context_t ctx, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                  uint64_t* d_out, int batch, void* stream);
int lsb_mul_plain_ringt(lsb_context_t ctx, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                        uint64_t* d_out, int batch, void* stream);
/* bind_gpu_cmp_sum / bind_gpu_cmpac_sum (mega_ag_executors_gpu.cu:294-408): out = sum_i ct_i x pt_i (+ acc).  d_cts[i] / d_pts[i]
 * are the batch blocks of term i (ct [2][level+1][N]; pt [level+1][N] NTT domain, or [1][N] mod t when ringt); d_acc may be NULL. */
int lsb_mac(lsb_context_t ctx, int level, int nterms, const uint64_t* const* d_cts, const uint64_t* const* d_pts, int ringt,
            const uint64_t* d_acc, uint64_t* d_out, int batch, void* stream);
/* Fused chains the runner substitutes for mult -> relin (-> rescale) node sequences
 * (frontend mult_relin, custom_task.py:1286; benchmark graphs examples/benchmark_gpu/benchmark_gpu.cpp:26-78). */
int lsb_mult_relin(lsb_context_t ctx, int level, const uint64_t* d_a, const uint64_t* d_b, const uint64_t* d_rlk,
                   int key_level, uint64_t* d_out2, int batch, void* stream);
int lsb_mult_relin_rescale(lsb_context_t ctx, int level, const uint64_t* d_a, const uint64_t* d_b,
                           const uint64_t* d_rlk, int key_level, uint64_t* d_out2, int batch, void* stream);

/* ---- CKKS bootstrapping (the `bootstrap` node: mega_ag_executors_gpu.cu:410-426, parameters gpu_wrapper.cu:106-116 /
 * frontend/custom_task.py:424-463; algorithm: lattisense_b200/csrc/lsb_bootstrap.h) ----
 * iparams[11] = {log_n, log_slots, cts_start_level, cts_depth, eval_mod_start_level, eval_mod_k, sine_deg, double_angle,
 *                stc_start_level, stc_depth, output_level};  dparams[4] = {scale, cts_bsgs_ratio, message_ratio, stc_bsgs_ratio}.
 * The plaintext matrices are encoded on the first run and stay on the device with the handle. */
typedef struct lsb_bootstrap_st* lsb_bootstrap_t;
lsb_bootstrap_t lsb_bootstrap_create(lsb_context_t ctx, const int* iparams, const double* dparams);
void lsb_bootstrap_destroy(lsb_bootstrap_t h);
/* rotation steps (mod N/2) whose Galois keys a run needs (besides the row key 2N-1); returns the count */
int lsb_bootstrap_rotations(lsb_bootstrap_t h, int* steps, int capacity);
/* d_ct0: [batch][2][1][N] level-0 ciphertexts; keys as the node receives them (custom_task.py:1961-1996): rlk, n_glk Galois
 * keys with their elements (all at key_level), swk_dts (level 0), swk_std (key_level); d_out: [batch][2][out_level+1][N].
 * stats5 (optional): rotations, key switches (all kinds), ct x ct, rescales, ct x pt. */
int lsb_bootstrap_run(lsb_bootstrap_t h, const uint64_t* d_ct0, const uint64_t* d_rlk, int key_level, const uint64_t* galois_elements,
                      const uint64_t* const* d_glk, int n_glk, const uint64_t* d_swk_dts, const uint64_t* d_swk_std, uint64_t* d_out,
                      int batch, void* stream, long long* stats5);

/* Launch accounting (always on) and per-kernel-class CUDA-event timing (bench.py roofline pass).
 * kind: 0 NTT fwd, 1 NTT inv, 2 basis extension, 3 key-switch inner product, 4 other limb-wise kernels. */
long long lsb_launch_count(lsb_context_t ctx);
void lsb_profile_enable(lsb_context_t ctx, int on);
int lsb_profile_read(lsb_context_t ctx, int kind, double* ms, long long* kernels, double* algorithmic_bytes);

/* ---- multi-GPU: limb-sharded hybrid key-switch (no reference counterpart -- the reference GPU
 * path is single-device, mega_ag_runners/gpu/gpu_wrapper.cu:141-391) ----
 * World of `world` ranks, one context per GPU.  Rank r owns Q limbs {i : i % world == r} and P limbs
 * {j : j % world == r} of every polynomial and of the evaluation key.  Buffers of one rank (u64 words, N = 2^log_n):
 *   c_own_ntt  [nq_own][N]          owned Q limbs of the polynomial to switch, NTT domain
 *   coeff_own  [cq][N]              phase-1 output (cq = ceil((level+1)/world); rows >= nq_own unused)
 *   coeff_all  [world][cq][N]       all-gather of coeff_own over the ranks (caller: NCCL)
 *   evk_own    [beta][2][nq_own+np_own][N]   owned limbs of the key (Q limbs first, then P limbs)
 *   acc_own    [2][nq_own+np_own][N]
 *   accP_own   [2][cp][N]           owned P limbs of the accumulators, coefficient domain (cp = ceil(k/world))
 *   accP_all   [world][2][cp][N]    all-gather of accP_own
 *   addend_own / out_own [2][nq_own][N]      out = ModDown(acc) (+ addend), owned Q limbs, NTT domain
 * Every limb produced equals the same limb of the single-GPU lsb_relin / lsb_rotate key-switch bit for bit. */
int lsb_ks_shard_layout(lsb_context_t ctx, int level, int world, int rank, int* nq_own, int* np_own, int* cq, int* cp,
                        int* own_q, int* own_p);
int lsb_ks_shard_begin(lsb_context_t ctx, int level, int world, int rank, const uint64_t* d_c_own_ntt,
                       uint64_t* d_coeff_own, int batch, void* stream);
int lsb_ks_shard_inner(lsb_context_t ctx, int level, int world, int rank, const uint64_t* d_coeff_all,
                       const uint64_t* d_c_own_ntt, const uint64_t* d_evk_own, uint64_t* d_acc_own, uint64_t* d_accP_own,
                       int batch, void* stream);
/* Fused compute + collective: the same phases with the all-gather done INSIDE the INTT kernel that produces the
 * limbs.  peer_*[r] is rank r's gathered buffer (coeff_all / accP_all) as mapped into THIS process (this rank's own
 * buffer at index `rank`; e.g. torch.distributed._symmetric_memory buffer_ptrs).  The kernel stores each output
 * word into all `world` buffers over NVLink peer memory; the caller only places a cross-rank barrier (no NCCL
 * collective) before the next phase reads the gathered buffer, and one before the buffers are overwritten again. */
int lsb_ks_shard_begin_p2p(lsb_context_t ctx, int level, int world, int rank, const uint64_t* d_c_own_ntt,
                           uint64_t* const* peer_coeff_all, void* stream);
int lsb_ks_shard_inner_p2p(lsb_context_t ctx, int level, int world, int rank, const uint64_t* d_coeff_all,
                           const uint64_t* d_c_own_ntt, const uint64_t* d_evk_own, uint64_t* d_acc_own,
                           uint64_t* const* peer_accP_all, void* stream);
int lsb_ks_shard_end(lsb_context_t ctx, int level, int world, int rank, const uint64_t* d_accP_all,
                     const uint64_t* d_acc_own, const uint64_t* d_addend_own /* may be NULL */, uint64_t* d_out_own,
                     int batch, void* stream);

/* ---- the same key-switch driven entirely from C/C++: one handle per rank owns the NCCL communicator, the exchange
 * buffers and a CUDA graph of the whole sequence (phase 1, all-gather, phase 2, all-gather, phase 3), batched over
 * `batch` ciphertexts.  NCCL is loaded at run time (libnccl.so.2).  Rank 0 calls lsb_nccl_unique_id and distributes the
 * 128 bytes to the other ranks by any means (torch.distributed broadcast, MPI, a file); every rank then creates its
 * handle with the same id (collective call).  Buffers per rank, densely packed per batch item:
 *   d_c_own_ntt [batch][nq_own][N], d_evk_own [beta][2][nq_own+np_own][N] (shared by the batch),
 *   d_addend_own (may be NULL) / d_out_own [batch][2][nq_own][N].
 * Every limb equals the single-GPU key-switch's bit for bit.  world = 1 degenerates to the single-GPU key-switch. */
typedef struct lsb_ks_shard_st* lsb_ks_shard_t;
int lsb_nccl_unique_id(void* id128);
lsb_ks_shard_t lsb_ks_shard_create(lsb_context_t ctx, int level, int world, int rank, const void* nccl_unique_id128, int max_batch);
void lsb_ks_shard_destroy(lsb_ks_shard_t h);
int lsb_ks_shard_run(lsb_ks_shard_t h, const uint64_t* d_c_own_ntt, const uint64_t* d_evk_own, const uint64_t* d_addend_own,
                     uint64_t* d_out_own, int batch, void* stream);
long long lsb_ks_shard_all_gather_bytes(lsb_ks_shard_t h); /* bytes this rank contributed to the two all-gathers of the last run */
const char* lsb_ks_shard_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LSB200_H */
