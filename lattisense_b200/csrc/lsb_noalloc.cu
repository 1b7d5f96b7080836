// lsb_mod_raise_static: CKKS ModRaise with NO device allocation -- the same kernel sequence as op_mod_raise
// (lsb_ops.cu), with the INTT's intermediate in a file-scope device array instead of the context's workspace.
//
// This entry point has no counterpart in the reference; it was added for
// callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The array is part of the library image: every GPU that loads the library holds LSB_STATIC_SCRATCH_WORDS * 8 bytes
// of it until the process exits.  On the host side the NAME `g_static_scratch` is only a stand-in for the symbol, not
// a device address; the address a kernel can use comes from cudaGetSymbolAddress, asked once per device with that
// device current (below), in this translation unit, which is the one that defines the array.
//
// Every call on a device shares the one array and the array keeps what the last call left in it.  ModRaise writes
// each word of the part it uses (the INTT's output) before it reads it, so nothing has to start at zero and nothing
// is cleared.  Two calls whose kernels overlap on the GPU (different streams without an ordering between them, or
// two host threads on two streams) RACE on the array: keep calls on one device ordered on the GPU.
#include <mutex>
#include <string>

#include "../../include/lsb200.h"
#include "lsb_ops.h"

using namespace lsb;

struct lsb_context_st;                              // lsb_capi.cu
Context& lsb_context_ref(lsb_context_t h);          // lsb_capi.cu
void lsb_set_last_error(const std::string& what);   // lsb_capi.cu: what lsb_last_error() returns on this thread

static __device__ __align__(256) unsigned long long g_static_scratch[LSB_STATIC_SCRATCH_WORDS];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mu;                           // guards g_addr and orders the host side of the calls
u64* g_addr[kMaxDevices] = {nullptr};      // per device: the array's DEVICE address

RowMap seq_rows(int n, int period) {  // rows 0..n-1 in place, modulus index r mod period (as rowmap_seq, lsb_ops.cu)
    RowMap m;
    m.n = n;
    for (int r = 0; r < n; r++) {
        m.in_slot[r] = (uint8_t)r;
        m.out_slot[r] = (uint8_t)r;
        m.mod[r] = (uint8_t)(r % period);
    }
    return m;
}

// once per device, with that device current
u64* static_scratch_on(int device) {
    LSB_CHECK(device >= 0 && device < kMaxDevices, "lsb_mod_raise_static: device index out of range");
    if (!g_addr[device]) {
        LSB_CUDA(cudaSetDevice(device));
        void* p = nullptr;
        cudaError_t err = cudaGetSymbolAddress(&p, g_static_scratch);
        if (err != cudaSuccess || !p)
            throw std::runtime_error(std::string("lsb_mod_raise_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(err));
        g_addr[device] = (u64*)p;
    }
    return g_addr[device];
}

}  // namespace

extern "C" int lsb_mod_raise_static(lsb_context_t h, int top_level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                                    void* stream) {
    try {
        lsb_set_last_error("");
        Context& c = lsb_context_ref(h);
        LSB_CHECK(c.scheme == SCHEME_CKKS, "lsb_mod_raise_static: mod_raise is a CKKS operation");
        LSB_CHECK(d_in && d_out, "lsb_mod_raise_static: null ciphertext pointer");
        LSB_CHECK(batch >= 1 && npoly >= 1, "lsb_mod_raise_static: batch and npoly must be positive");
        LSB_CHECK(top_level >= 1 && top_level < c.nq && npoly * (top_level + 1) <= kMaxRows, "lsb_mod_raise_static: bad level");
        const size_t item = (size_t)npoly * c.N;  // words of scratch per ciphertext: the level-0 limbs, coefficient domain
        if (item * (size_t)batch > (size_t)LSB_STATIC_SCRATCH_WORDS)
            throw std::runtime_error("lsb_mod_raise_static: npoly * N * batch = " + std::to_string(item * (size_t)batch) +
                                     " words of scratch exceed the static array's LSB_STATIC_SCRATCH_WORDS = " +
                                     std::to_string((size_t)LSB_STATIC_SCRATCH_WORDS) + "; split the batch or call lsb_mod_raise");
        std::lock_guard<std::mutex> lk(g_mu);
        u64* scratch = static_scratch_on(c.device);
        LSB_CUDA(cudaSetDevice(c.device));
        cudaStream_t s = (cudaStream_t)stream;
        const int l = top_level + 1;
        BatchPtr in = bp_strided(d_in, c.ct_words(0, npoly));
        BatchPtr out = bp_strided(d_out, c.ct_words(top_level, npoly));
        BatchPtr last = bp_strided(scratch, item);
        launch_ntt(c.dt, in, last, seq_rows(npoly, 1), batch, true, s);
        launch_rescale_lift(c.dt, last, 1, out, npoly, top_level, batch, s, true);
        launch_ntt(c.dt, out, out, seq_rows(npoly * l, l), batch, false, s);
        return 0;
    } catch (const std::exception& ex) {
        lsb_set_last_error(ex.what());
        return 1;
    } catch (...) {
        lsb_set_last_error("lsb_mod_raise_static: unknown exception");
        return 1;
    }
}
