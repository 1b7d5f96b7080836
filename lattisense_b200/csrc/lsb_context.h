// RNS context of one parameter set on one GPU: moduli, twiddle tables, cached conversion tables.
// Replaces init_gpu_context (mega_ag_runners/gpu/gpu_wrapper.cu:53-138), which the reference
// rebuilds on every run(); here it is built once per (task, device) and reused.
#pragma once
#include <mutex>
#include <unordered_map>
#include <memory>
#include "lsb_kernels.h"
#include "lsb_shard.h"

namespace lsb {

enum Scheme { SCHEME_BFV = 0, SCHEME_CKKS = 1 };

// Grow-only device scratch; one per stream that runs ops.
class Workspace {
public:
    explicit Workspace(int device) : device_(device) {}
    ~Workspace();
    Workspace(const Workspace&) = delete;
    Workspace& operator=(const Workspace&) = delete;
    // returns a buffer of at least `words` u64, valid until the next reserve() on this workspace
    u64* reserve(size_t words, cudaStream_t s);
    size_t capacity_words() const { return cap_; }

private:
    int device_;
    u64* buf_ = nullptr;
    size_t cap_ = 0;
};

struct Context {
    int device = 0;
    int scheme = SCHEME_CKKS;
    int logN = 0, N = 0;
    int nq = 0, np = 0, nqm = 0;
    u64 t = 0;
    std::vector<u64> moduli;      // Q..., P..., QMul...
    std::vector<int> level_qmul;  // BFV: index of the last auxiliary prime used at each level
    DeviceTables dt;
    Profiler prof;

    Context(int device, int scheme, int logN, const u64* q, int nq, const u64* p, int np, u64 t);
    ~Context();
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;

    int idxP(int j) const { return nq + j; }
    int idxQM(int j) const { return nq + np + j; }
    int beta(int level) const { return (level + 1 + np - 1) / np; }
    size_t ct_words(int level, int npoly) const { return (size_t)npoly * (level + 1) * N; }
    size_t evk_words(int key_level) const { return (size_t)beta(key_level) * 2 * (key_level + 1 + np) * N; }

    // ---- cached device-side tables (built on first use; thread-safe) ----
    int ks_digit_group(int level) const;                    // digits per launch group of the key-switch ModUp (>= beta: one group)
    const BexJob* modup_jobs(int level, int coeff_slot0);   // beta jobs; destination slots relative to the digit's group
    const BexJob* moddown_jobs(int level);                  // 2 jobs (one per polynomial)
    const FinishMap* moddown_finish(int level, int addmask, int out_stride);  // rows (p,i); addmask bit p: add c_p
    const FinishMap* rescale_finish(int level, int npoly);
    const FinishMap* finish_host(const FinishMap* dev);
    const BexMeta* bex_meta(const BexJob* dev);  // host description of a job array returned by the *_jobs builders  // host copy of a map returned by the two builders above
    const Tw* qlinv(int level);                             // [level]  q_level^-1 mod q_i
    const u32* permute_index(u64 gal);
    const BexJob* bfv_q_to_qmul_jobs(int level);            // 2 jobs: input ct -> extended ct
    const BexJob* bfv_quant1_jobs(int level);               // 3 jobs
    const BexJob* bfv_quant2_jobs(int level);               // 3 jobs
    const ScaleUpConsts* bfv_scale_up_consts(int level);

    Workspace& default_workspace() { return *ws0_; }
    // side stream + fork/join events of the two-stream batch pipeline (lsb_ops.cu: mult_relin_chain); created on first use
    struct SidePipe {
        cudaStream_t stream = nullptr;
        cudaEvent_t fork = nullptr, join = nullptr;
    };
    SidePipe& side_pipe();
    // limb-sharded key-switch tables (lsb_shard.cu), cached per (level, world, rank)
    const ShardPlan& shard_plan(int level, int world, int rank);
    const void* upload_raw(const void* host, size_t bytes) {
        std::lock_guard<std::mutex> lk(mu_);
        return to_device(host, bytes);
    }

private:
    void* to_device(const void* host, size_t bytes);
    BexJob make_bex(const std::vector<int>& src_mod, const std::vector<int>& dst_mod) const;
    std::mutex mu_;
    std::vector<void*> owned_;
    std::unordered_map<u64, const void*> cache_;
    std::unordered_map<const void*, FinishMap> fm_host_;
    std::mutex shard_mu_;
    std::unordered_map<u64, std::unique_ptr<ShardPlan>> shard_plans_;
    std::unordered_map<const void*, std::vector<BexMeta>> bex_meta_;
    const void* upload_jobs(const std::vector<BexJob>& jobs);
    std::unique_ptr<Workspace> ws0_;
    SidePipe side_;
};

}  // namespace lsb
