/*
 * lattisense_gpu_runner.h — the drop-in boundary of libgpu_mega_ag_runner.so.
 *
 * These are the entry points the reference's SDK and plug-ins bind for the GPU path
 * (paths relative to cipherflow-fhe/lattisense @ 88bf833):
 *   - cxx_sdk_v2/cxx_fhe_task_gpu.cpp:32-117 (FheTaskGpu) links them directly,
 *   - plug-in/lattigo/acc/gpu_runner.go:6 (cgo, -lgpu_mega_ag_runner),
 *   - plug-in/SEAL/acc/CMakeLists.txt:14.
 * Prototypes, argument meaning and error behaviour follow mega_ag_runners/wrapper.h:67-85 and the
 * implementation in mega_ag_runners/gpu/gpu_wrapper.cu:481-530; the argument and ciphertext/key
 * structs are byte-compatible restatements of mega_ag_runners/c_argument.h:24-46 and
 * abi/c_types.h:26-60.  The `void*` executor parameters are pointers to C++ `ExecutorFunc`
 * objects (see lattisense_mega_ag.hpp) which are copied during the call, exactly as in the reference.
 *
 * Errors: create/bind throw std::runtime_error like the reference (mega_ag.cpp:136-138); run_fhe_gpu_task
 * never throws: it returns non-zero when any node failed (the SDK turns that into a throw,
 * cxx_fhe_task_gpu.cpp:106-108) and lsb_gpu_task_last_error() holds the message.
 * There is no CPU fallback: without a CUDA device run_fhe_gpu_task fails.
 */
#ifndef LATTISENSE_GPU_RUNNER_H
#define LATTISENSE_GPU_RUNNER_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- mega_ag_runners/c_argument.h:24-46 ---- */
#ifndef LATTISENSE_C_ARGUMENT_DEFINED
#define LATTISENSE_C_ARGUMENT_DEFINED
typedef enum {
    TYPE_PLAINTEXT,
    TYPE_CIPHERTEXT,
    TYPE_RELIN_KEY,
    TYPE_GALOIS_KEY,
    TYPE_SWITCH_KEY,
    TYPE_CUSTOM,
} DataType;

typedef enum {
    ALGO_BFV,
    ALGO_CKKS,
} Algo;

typedef struct {
    const char* id; /* argument name from task_signature.json */
    DataType type;
    void* data; /* void*[size]: opaque frontend handles, only ever dereferenced by the bound executors */
    int level;
    int size;
} CArgument;
#endif

/* ---- abi/c_types.h:26-60: wire layout of ciphertexts, plaintexts and keys ---- */
#ifndef LATTISENSE_C_TYPES_DEFINED
#define LATTISENSE_C_TYPES_DEFINED
typedef struct {
    int n;          /* ring degree N */
    uint64_t* data; /* N residues in [0, q_i), malloc'd per limb (abi/c_structs.c:23-26) */
} CComponent;

typedef struct {
    int n_component; /* level+1 limbs (keys: level+1 Q limbs then k P limbs) */
    CComponent* components;
} CPolynomial;

typedef struct {
    int level;
    CPolynomial poly;
} CPlaintext;

typedef struct {
    int level;
    int degree; /* polys = degree+1 */
    CPolynomial* polys;
} CCiphertext;

typedef CCiphertext CPublicKey;

typedef struct {
    int n_public_key; /* beta = ceil((level+1)/k) digits, each a degree-1 "ciphertext" over Q u P */
    CPublicKey* public_keys;
} CKeySwitchKey;

typedef CKeySwitchKey CRelinKey;

typedef struct {
    int n_key_switch_key;
    uint64_t* galois_elements;
    CKeySwitchKey* key_switch_keys;
} CGaloisKey;
#endif

/* ---- abi/c_structs.h:22-44, abi/c_structs.c:23-98: allocation helpers (same malloc-per-limb ownership) ---- */
void alloc_component(CComponent* component, int n);
void alloc_plaintext(CPlaintext* pt, int level, int n);
void alloc_ciphertext(CCiphertext* ct, int degree, int level, int n);
void alloc_relin_key(CRelinKey* rlk, int n_public_key, int level, int n);
void set_galois_key_steps(CGaloisKey* glk, uint64_t* galois_elements, int n_galois_elements);
void free_plaintext(CPlaintext* pt);
void free_ciphertext(CCiphertext* ct);
void free_relin_key(CRelinKey* rlk);
void free_galois_key(CGaloisKey* gk);

/* ---- mega_ag_runners/wrapper.h:26-42, 67-85 ---- */
typedef struct fhe_task_handle_st* fhe_task_handle;
/* invoked from worker threads, (completed, total) compute nodes */
typedef void (*progress_callback_t)(int completed, int total, void* user_data);

/* Loads <project_path>/mega_ag.json (gpu_wrapper.cu:412-431). */
fhe_task_handle create_fhe_gpu_task(const char* project_path);
void release_fhe_gpu_task(fhe_task_handle handle);
/* abi_export_executor / abi_import_executor: ExecutorFunc* (gpu_wrapper.cu:492-497). */
void bind_gpu_task_abi_bridge_executors(fhe_task_handle handle, void* abi_export_executor, void* abi_import_executor);
/* executors[i]: ExecutorFunc* for custom node type custom_types[i] (gpu_wrapper.cu:499-510). */
void bind_gpu_task_custom_executors(fhe_task_handle handle, const char** custom_types, void** executors,
                                    uint64_t n_executors);
/* Runs the task graph on GPU `gpu_device`; synchronous; 0 on success (gpu_wrapper.cu:512-529).
 * Inputs are flattened in mega_ag.inputs order (keys after data: rlk, glk, swk_dts, swk_std;
 * cxx_sdk_v2/cxx_argument.h:178-260), outputs map positionally onto mega_ag.outputs. */
int run_fhe_gpu_task(fhe_task_handle handle, CArgument* input_args, uint64_t n_in_args, CArgument* output_args,
                     uint64_t n_out_args, progress_callback_t progress_cb, void* user_data, int gpu_device);

/* ---- extensions of this implementation (not in the reference; optional for callers) ---- */
/* Devices used by the NEXT run of this task: independent components of the DAG are sharded over
 * devices[0..n).  Default: the single `gpu_device` passed to run_fhe_gpu_task. */
void lsb_gpu_task_set_devices(fhe_task_handle handle, const int* devices, int n_devices);
/* The list is a request: a device is opened only per LSB_MIN_WEIGHT_PER_DEVICE (environment, default 800 = about 200 key
 * switches at n = 16384; 0 = always use every device) of work in the graph and never for more devices than the graph has
 * independent components, because each device costs a copy of the keys and a shorter launch train (cfg3 x 64 on B200s:
 * 11.7 ms on 4 devices, 15.5 ms on 8).  Returns how many of `n_devices` the next run would use (the first ones listed). */
int lsb_gpu_task_worthwhile_devices(fhe_task_handle handle, int n_devices);
/* Message of the last failed run_fhe_gpu_task on the calling thread ("" if none). */
const char* lsb_gpu_task_last_error(void);
/* Statistics of the last run: kernel launches, batched launch groups, nodes fused away. */
void lsb_gpu_task_last_stats(fhe_task_handle handle, long long* kernel_launches, long long* batched_groups,
                             long long* fused_nodes);

/* Device-resident key cache across run() calls (the reference uploads every key on every run,
 * mega_ag_runners/gpu/gpu_abi_bridge_executors.h:213-310): with the cache on, a key HANDLE (the opaque pointer in
 * CArgument.data) that was passed to an earlier run of this task is neither exported nor uploaded again -- the caller
 * promises that the handle still holds the same key.  Off by default (or LSB_KEY_CACHE=1 in the environment);
 * lsb_gpu_task_invalidate_keys drops the cached keys (call it after re-generating keys in place). */
void lsb_gpu_task_set_key_cache(fhe_task_handle handle, int on);
void lsb_gpu_task_invalidate_keys(fhe_task_handle handle);
/* Transfer accounting of the last run: bytes copied host->device and device->host, the part of the former that was
 * read straight out of the pinned staging memory lent to the EXPORT executor (no second host copy), and key bytes
 * not uploaded thanks to the key cache. */
void lsb_gpu_task_transfer_stats(fhe_task_handle handle, long long* h2d_bytes, long long* d2h_bytes, long long* zero_copy_bytes,
                                 long long* key_bytes_cached);

/* Diagnostics of the execution plan built at create_fhe_gpu_task (used by this repository's host-side tests):
 * out8 = {compute nodes incl. bridge nodes, data nodes, host units, LOAD units, STORE units, GPU units, nodes fused away,
 * independent DAG components}; returns 0. */
int lsb_gpu_task_describe(fhe_task_handle handle, long long* out8);
/* component -> device slot for a run over `n_slots` devices (longest-processing-time-first); writes at most `capacity`
 * entries, returns the number of components. */
int lsb_gpu_task_shard_preview(fhe_task_handle handle, int n_slots, int* slot_of_component, int capacity);

/* Copy helper for plug-in EXPORT / IMPORT executors: memcpy semantics, streaming (non-temporal) stores for large aligned
 * blocks.  Limbs handed out by alloc_component while an EXPORT executor runs are pinned staging memory that the CPU never
 * reads again; filling them with this instead of memcpy removes the read-for-ownership traffic (a third of the copy). */
void lsb_host_copy(void* dst, const void* src, size_t bytes);
/* Host-side time of the last run in ms: [0] wall time of run_fhe_gpu_task, then SUMS over the worker threads of
 * [1] EXPORT executors, [2] staging copies, [3] waiting for pinned-ring space, [4] waiting for a result's D2H copy,
 * [5] IMPORT executors. */
void lsb_gpu_task_host_profile(fhe_task_handle handle, double* ms6);

#ifdef __cplusplus
}
#endif
#endif /* LATTISENSE_GPU_RUNNER_H */
