// Launch accounting and optional per-kernel-class CUDA-event timing (bench.py's roofline pass).
// Counting is always on (one relaxed atomic add per launch); event timing only when enabled.
#pragma once
#include <atomic>
#include <mutex>
#include <vector>
#include <cuda_runtime.h>

namespace lsb {

enum ProfKind { PK_NTT_FWD = 0, PK_NTT_INV, PK_BASIS_EXT, PK_KS_INNER, PK_ELEMWISE, PK_COUNT };

struct Profiler {
    std::atomic<long long> launches{0};
    std::atomic<bool> enabled{false};
    struct Rec {
        int kind;
        cudaEvent_t e0, e1;
        double alg_bytes;
        int kernels;
    };
    std::mutex mu;
    std::vector<Rec> recs;
    void reset();
    // synchronises the recorded events and sums them
    void read(int kind, double* ms, long long* kernels, double* alg_bytes);
    ~Profiler() { reset(); }
};

struct ProfScope {
    Profiler* p;
    cudaStream_t s;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int kind, kernels;
    double bytes;
    ProfScope(Profiler* prof, int kind_, cudaStream_t s_, double alg_bytes, int kernels_ = 1)
        : p(prof), s(s_), kind(kind_), kernels(kernels_), bytes(alg_bytes) {
        if (!p) return;
        p->launches.fetch_add(kernels, std::memory_order_relaxed);
        if (p->enabled.load(std::memory_order_relaxed)) {
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            cudaEventRecord(e0, s);
        }
    }
    ~ProfScope() {
        if (!e0) return;
        cudaEventRecord(e1, s);
        std::lock_guard<std::mutex> lk(p->mu);
        p->recs.push_back({kind, e0, e1, bytes, kernels});
    }
};

}  // namespace lsb
