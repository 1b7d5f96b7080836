// Staging copy pageable -> pinned with non-temporal stores.
//
// The LOAD path copies every limb once into the pinned ring the H2D DMA reads from.  The destination is never
// read by the CPU again, so a regular memcpy wastes a third of the host memory traffic on read-for-ownership of
// the destination lines; streaming stores skip that (measured on the B200 host, tools/host_bw.cu: 15 threads,
// 128 KiB blocks: 60 GB/s memcpy vs 80 GB/s streaming; with PCIe busy 39 vs 62 GB/s).
#include <immintrin.h>
#include <stdint.h>
#include <string.h>

namespace lsr {

__attribute__((target("avx2"))) static void stream_copy_avx2(uint8_t* dst, const uint8_t* src, size_t bytes) {
    const __m256i* s = (const __m256i*)src;
    __m256i* d = (__m256i*)dst;
    const size_t n = bytes / 32;
    size_t i = 0;
    for (; i + 4 <= n; i += 4) {
        const __m256i a = _mm256_loadu_si256(s + i), b = _mm256_loadu_si256(s + i + 1);
        const __m256i c = _mm256_loadu_si256(s + i + 2), e = _mm256_loadu_si256(s + i + 3);
        _mm256_stream_si256(d + i, a);
        _mm256_stream_si256(d + i + 1, b);
        _mm256_stream_si256(d + i + 2, c);
        _mm256_stream_si256(d + i + 3, e);
    }
    for (; i < n; i++) _mm256_stream_si256(d + i, _mm256_loadu_si256(s + i));
    _mm_sfence();
    if (bytes & 31) memcpy(dst + n * 32, src + n * 32, bytes & 31);
}

void lsr_stream_copy(void* dst, const void* src, size_t bytes) {
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2 && ((uintptr_t)dst & 31) == 0 && bytes >= 4096) stream_copy_avx2((uint8_t*)dst, (const uint8_t*)src, bytes);
    else memcpy(dst, src, bytes);
}

}  // namespace lsr

// Public copy helper for plug-in EXPORT / IMPORT executors (declared in include/lattisense_gpu_runner.h): limbs handed
// out by alloc_component during an EXPORT are pinned staging memory the CPU never reads again, so filling them with
// streaming stores saves the read-for-ownership third of the host memory traffic -- and host DRAM bandwidth, not PCIe
// or the GPU, is what bounds run_fhe_gpu_task end to end (profiles/r02_e2e_host_path.txt).
extern "C" void lsb_host_copy(void* dst, const void* src, size_t bytes) { lsr::lsr_stream_copy(dst, src, bytes); }
