// Shared host/device declarations for the B200 executor core.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <stdexcept>
#include <string>
#include <vector>
#include <map>
#include <memory>
#include <cuda_runtime.h>
#include "lsb_math.h"
#include "lsb_ntt_arith.h"

namespace lsb {

struct CudaError : std::runtime_error {
    explicit CudaError(const std::string& s) : std::runtime_error(s) {}
};
#define LSB_CUDA(call)                                                                                  \
    do {                                                                                                \
        cudaError_t _e = (call);                                                                        \
        if (_e != cudaSuccess)                                                                          \
            throw ::lsb::CudaError(std::string(#call) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                                   ":" + std::to_string(__LINE__) + ")");                               \
    } while (0)
#define LSB_CHECK(cond, msg)                                   \
    do {                                                       \
        if (!(cond)) throw std::runtime_error(std::string(msg)); \
    } while (0)

// Where a batch item's block of limbs lives: either table[b] (independently allocated
// ciphertexts, as the mega_ag runner holds them) or base + b*stride (runner-owned scratch).
struct BatchPtr {
    u64* const* table;
    u64* base;
    size_t stride;  // words
#if defined(__CUDACC__)
    __device__ __forceinline__ u64* at(int b) const { return table ? table[b] : base + (size_t)b * stride; }
#endif
};
inline BatchPtr bp_strided(u64* base, size_t stride) { return BatchPtr{nullptr, base, stride}; }
inline BatchPtr bp_strided(const u64* base, size_t stride) { return BatchPtr{nullptr, const_cast<u64*>(base), stride}; }
inline BatchPtr bp_table(u64* const* t) { return BatchPtr{t, nullptr, 0}; }
inline BatchPtr bp_offset(BatchPtr p, size_t words) {
    // only valid for strided pointers; table pointers carry their offset in the row map
    p.base += words;
    return p;
}

// Rows (limbs) of one batch item a kernel works on: slot = limb offset (in units of N words)
// inside the item's input / output block, mod = index into the context's modulus table.
// Slots and modulus indices are uint8_t: a launch addresses at most kMaxRows = 192 limbs of one item and slot numbers stay
// below 256 -- the largest user is the cfg5 key-switch, beta * (l + k) = 5 * 30 = 150 extended limbs (every builder checks
// against kMaxRows; a parameter set with more than 255 limbs per item would need uint16_t here).
constexpr int kMaxRows = 192;
struct RowMap {
    int n;
    uint8_t in_slot[kMaxRows];
    uint8_t out_slot[kMaxRows];
    uint8_t mod[kMaxRows];
};

// Per-modulus constants resident on the device.
struct ModConst {
    u64 q;
    u64 qinv64;  // floor((2^64-1)/q)
    BarrettQ bq;
    Tw ninv;    // N^-1
    Tw w_ninv;  // psi^-1_br[1] * N^-1
    // FP64 butterfly policy (lsb_ntt_arith.h); f64_ok == 0 for primes that are too wide
    int f64_ok, f64_recentre;
    double qd, qinvd;
    TwD ninv_d, w_ninv_d;
};

}  // namespace lsb
