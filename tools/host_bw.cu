// Host-side copy and PCIe bandwidth probe for the runner's LOAD/STORE path (development aid).
//   nvcc -O3 -std=c++17 -Xcompiler -mavx2,-pthread tools/host_bw.cu -o tools/_build/host_bw
// Prints: multi-threaded 128 KiB-block copy rates pageable->pinned (glibc memcpy vs non-temporal stores),
// H2D / D2H rates from pinned memory alone and concurrently, and the copy rate while PCIe is busy.
#include <cuda_runtime.h>
#include <immintrin.h>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static void nt_copy(void* dst, const void* src, size_t bytes) {
    const __m256i* s = (const __m256i*)src;
    __m256i* d = (__m256i*)dst;
    for (size_t i = 0; i < bytes / 32; i += 4) {
        __m256i a = _mm256_loadu_si256(s + i), b = _mm256_loadu_si256(s + i + 1), c = _mm256_loadu_si256(s + i + 2),
                e = _mm256_loadu_si256(s + i + 3);
        _mm256_stream_si256(d + i, a);
        _mm256_stream_si256(d + i + 1, b);
        _mm256_stream_si256(d + i + 2, c);
        _mm256_stream_si256(d + i + 3, e);
    }
    _mm_sfence();
}

int main() {
    const size_t BLK = 128 << 10, NBLK = 8192;  // 1 GiB of pageable blocks
    std::vector<uint8_t*> blocks(NBLK);
    for (auto& b : blocks) {
        if (posix_memalign((void**)&b, 64, BLK)) return 1;
        memset(b, 1, BLK);
    }
    uint8_t* pinned;
    cudaHostAlloc((void**)&pinned, BLK * NBLK, cudaHostAllocDefault);
    memset(pinned, 2, BLK * NBLK);
    uint8_t* dev;
    cudaMalloc((void**)&dev, BLK * NBLK);
    cudaStream_t s1, s2;
    cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    setvbuf(stdout, nullptr, _IONBF, 0);
    printf("hw threads %u\n", std::thread::hardware_concurrency());

    auto run_copy = [&](int T, int mode, bool to_pinned) {
        std::atomic<size_t> next{0};
        double t0 = now();
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++)
            th.emplace_back([&] {
                for (;;) {
                    size_t i = next.fetch_add(1);
                    if (i >= NBLK) break;
                    uint8_t* p = pinned + i * BLK;
                    if (to_pinned) (mode ? nt_copy(p, blocks[i], BLK) : (void)memcpy(p, blocks[i], BLK));
                    else (mode ? nt_copy(blocks[i], p, BLK) : (void)memcpy(blocks[i], p, BLK));
                }
            });
        for (auto& x : th) x.join();
        return (double)BLK * NBLK / (now() - t0) / 1e9;
    };
    for (int T : {1, 2, 4, 8, 12, 15, 16, 24, 32}) {
        run_copy(T, 0, true);
        printf("T=%2d pageable->pinned memcpy %6.1f GB/s  nt %6.1f GB/s | pinned->pageable memcpy %6.1f  nt %6.1f\n", T,
               run_copy(T, 0, true), run_copy(T, 1, true), run_copy(T, 0, false), run_copy(T, 1, false));
    }
    // malloc+free cost of a 128 KiB limb (the ABI allocates one per limb)
    {
        double t0 = now();
        for (int i = 0; i < 20000; i++) {
            void* p = malloc(BLK);
            ((volatile char*)p)[0] = 1;
            free(p);
        }
        printf("malloc+touch+free 128KiB: %.2f us\n", (now() - t0) / 20000 * 1e6);
    }
    auto pcie = [&](bool h2d, bool d2h) {
        const size_t half = BLK * NBLK / 2;
        cudaDeviceSynchronize();
        double t0 = now();
        for (int r = 0; r < 4; r++) {
            if (h2d) cudaMemcpyAsync(dev, pinned, half, cudaMemcpyHostToDevice, s1);
            if (d2h) cudaMemcpyAsync(pinned + half, dev + half, half, cudaMemcpyDeviceToHost, s2);
        }
        cudaDeviceSynchronize();
        return 4.0 * half / (now() - t0) / 1e9;
    };
    pcie(true, true);
    printf("H2D alone %.1f GB/s, D2H alone %.1f GB/s, concurrent %.1f GB/s each\n", pcie(true, false), pcie(false, true),
           pcie(true, true));
    // 1 MiB copies (one ciphertext each) back to back
    {
        cudaDeviceSynchronize();
        double t0 = now();
        for (size_t i = 0; i < 1024; i++) cudaMemcpyAsync(dev + (i << 20), pinned + (i << 20), 1 << 20, cudaMemcpyHostToDevice, s1);
        cudaDeviceSynchronize();
        printf("H2D 1024 x 1 MiB: %.1f GB/s\n", 1024.0 * (1 << 20) / (now() - t0) / 1e9);
    }
    // host copies while PCIe is busy in both directions
    {
        std::atomic<bool> stop{false};
        std::thread bg([&] {
            while (!stop.load()) pcie(true, true);
        });
        double r = run_copy(15, 1, true);
        double r2 = run_copy(15, 0, true);
        stop = true;
        bg.join();
        printf("with PCIe busy: T=15 nt %.1f GB/s memcpy %.1f GB/s\n", r, r2);
    }
    return 0;
}
