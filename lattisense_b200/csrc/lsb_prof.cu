#include "lsb_prof.h"

namespace lsb {

void Profiler::reset() {
    std::lock_guard<std::mutex> lk(mu);
    for (auto& r : recs) {
        cudaEventDestroy(r.e0);
        cudaEventDestroy(r.e1);
    }
    recs.clear();
}

void Profiler::read(int kind, double* ms, long long* kernels, double* alg_bytes) {
    std::lock_guard<std::mutex> lk(mu);
    double t = 0, b = 0;
    long long n = 0;
    for (auto& r : recs) {
        if (r.kind != kind) continue;
        cudaEventSynchronize(r.e1);
        float f = 0;
        cudaEventElapsedTime(&f, r.e0, r.e1);
        t += f;
        b += r.alg_bytes;
        n += r.kernels;
    }
    *ms = t;
    *kernels = n;
    *alg_bytes = b;
}

}  // namespace lsb
