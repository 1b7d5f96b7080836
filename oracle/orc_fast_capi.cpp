// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.hpp header).  C ABI of the optimised CPU baseline (orc_fast.hpp):
// used by bench.py's reference arm / cpu_baseline leg and by tests/test_oracle_cpu.py (equality with the simple oracle).
#include "orc_fast.hpp"

using namespace orc;

struct FastHandle {
    std::unique_ptr<Context> base;
    std::unique_ptr<orcf::FContext> fc;
};

extern "C" {

void* orcf_create(int scheme, int logN, const u64* q, int nq, const u64* p, int np, u64 t) {
    try {
        auto h = new FastHandle;
        h->base = make_context(scheme, logN, q, nq, p, np, t);
        h->fc = orcf::make_fcontext(*h->base);
        return h;
    } catch (const std::exception&) {
        return nullptr;
    }
}
void orcf_destroy(void* h) { delete (FastHandle*)h; }
void orcf_ntt(void* h, int midx, u64* a, int inverse) {
    orcf::FContext& c = *((FastHandle*)h)->fc;
    if (inverse) orcf::ntt_inverse(c.mods.at(midx), c.N, a);
    else orcf::ntt_forward(c.mods.at(midx), c.N, a);
}
void orcf_ckks_mult_relin_rescale(void* h, int level, const u64* a, const u64* b, const u64* rlk, int key_level, u64* out) {
    orcf::ckks_mult_relin_rescale(*((FastHandle*)h)->fc, level, a, b, rlk, key_level, out);
}
void orcf_ckks_mult_relin(void* h, int level, const u64* a, const u64* b, const u64* rlk, int key_level, u64* out) {
    orcf::FContext& c = *((FastHandle*)h)->fc;
    orcf::ckks_mult(c, level, a, b, c.ct3.data());
    orcf::ckks_relin(c, level, c.ct3.data(), rlk, key_level, out);
}
void orcf_ckks_rotate(void* h, int level, const u64* ct, const u64* glk, int key_level, u64 gal, u64* out) {
    FastHandle* fh = (FastHandle*)h;
    auto idx = permute_ntt_index(fh->base->logN, gal);
    orcf::ckks_rotate(*fh->fc, level, ct, glk, key_level, idx.data(), out);
}
// repeat count inside the library so the timed loop has no Python in it
double orcf_time_mult_relin_rescale(void* h, int level, const u64* a, const u64* b, const u64* rlk, int key_level, u64* out,
                                    int reps) {
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < reps; r++) orcf::ckks_mult_relin_rescale(*((FastHandle*)h)->fc, level, a, b, rlk, key_level, out);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

}  // extern "C"
