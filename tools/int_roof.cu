// Integer-pipe microbenchmark for sm_100a: issue rates of the instruction classes a 64-bit Shoup
// butterfly is made of (IMAD, IMAD.WIDE.U32, IMAD.HI.U32, IADD3, 64-bit mul.lo / mul.hi, whole butterfly).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/int_roof tools/int_roof.cu
// Prints lanes/clk/SM for each class and butterflies/clk/SM; used for the integer co-roof in DESIGN.md.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint64_t u64;
typedef uint32_t u32;

constexpr int ITERS = 4096;
constexpr int ILP = 8;

template <int KIND>
__global__ void __launch_bounds__(1024) k(u64* out, u64 seed, u64 q) {
    u64 x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = seed + threadIdx.x * 977 + i * 131;
    u32 a32[ILP], b32 = (u32)seed | 1, c32 = (u32)(seed >> 7) | 3;
#pragma unroll
    for (int i = 0; i < ILP; i++) a32[i] = (u32)x[i];
    const u64 w = seed * 0x9E3779B97F4A7C15ull % q, ws = (u64)(((unsigned __int128)w << 64) / q);
    const u64 q2 = 2 * q, q4 = 4 * q;
    const u32 w1p = (u32)(ws >> 32);
    const double dw0s = (double)(u32)ws * 2.3283064365386963e-10, dw1s = (double)w1p * 2.3283064365386963e-10;  // * 2^-32
    double d[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) d[i] = (double)(x[i] & 0xffffffffffull);
    const double dq = (double)(q >> 14), dw = (double)(w >> 14), dwq = dw / dq;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (KIND == 0) {  // IMAD 32-bit
                a32[i] = a32[i] * b32 + c32;
            } else if (KIND == 1) {  // IMAD.WIDE.U32
                x[i] = (u64)(u32)x[i] * b32 + x[i];
            } else if (KIND == 2) {  // IMAD.HI.U32
                a32[i] = __umulhi(a32[i], b32) + c32;
            } else if (KIND == 3) {  // IADD3 / LOP3 (alu pipe)
                a32[i] = (a32[i] + b32) ^ c32;
            } else if (KIND == 4) {  // mul.lo.u64
                x[i] = x[i] * w + ws;
            } else if (KIND == 5) {  // mul.hi.u64
                x[i] = __umul64hi(x[i], ws) + w;
            } else if (KIND == 6) {  // Shoup lazy multiply
                u64 h = __umul64hi(x[i], ws);
                x[i] = x[i] * w - h * q;
            } else if (KIND == 7) {  // Harvey butterfly on pairs (ILP/2 butterflies)
                if (i & 1) {
                    u64 X = x[i - 1], Y = x[i];
                    X = X >= q2 ? X - q2 : X;
                    u64 h = __umul64hi(Y, ws);
                    u64 T = Y * w - h * q;
                    x[i - 1] = X + T;
                    x[i] = X - T + q2;
                }
            } else if (KIND == 9) {  // DFMA
                d[i] = fma(d[i], dw, dq);
            } else if (KIND == 10) {  // FP64 lazy butterfly (q < 2^48): 8 FP64-pipe ops
                if (i & 1) {
                    double X = d[i - 1], Y = d[i];
                    double h = Y * dw;
                    double l = fma(Y, dw, -h);
                    double qm = fma(Y, dwq, 6755399441055744.0);
                    double qf = qm - 6755399441055744.0;
                    double r = fma(-qf, dq, h) + l;
                    d[i - 1] = X + r;
                    d[i] = X - r;
                }
            } else if (KIND == 11) {  // FP64 butterfly interleaved with an integer Shoup butterfly (pipe overlap)
                if (i & 1) {
                    double X = d[i - 1], Y = d[i];
                    double h = Y * dw;
                    double l = fma(Y, dw, -h);
                    double qm = fma(Y, dwq, 6755399441055744.0);
                    double qf = qm - 6755399441055744.0;
                    double r = fma(-qf, dq, h) + l;
                    d[i - 1] = X + r;
                    d[i] = X - r;
                    u64 XX = x[i - 1], YY = x[i];
                    u64 hh = __umul64hi(YY, ws);
                    u64 T = YY * w - hh * q;
                    x[i - 1] = XX + T;
                    x[i] = XX - T + q2;
                }
            } else if (KIND == 12) {  // hybrid Shoup multiply: the high product's cross terms on the FP64 pipe
                // y*w' / 2^64 = y1*w1' + (y1*w0' + y0*w1') / 2^32 + y0*w0' / 2^64: the first term exact (one IMAD.WIDE), the
                // middle one floor-rounded in two DFMA.RM on the 2^52 grid, the last dropped: qhat in [Qshoup - 2, Qshoup],
                // r = y*w - qhat*q in [0, 4q)
                const u32 y0 = (u32)x[i], y1 = (u32)(x[i] >> 32);
                const double dy1 = __hiloint2double(0x43300000, (int)y1) - 4503599627370496.0;
                const double dy0 = __hiloint2double(0x43300000, (int)y0) - 4503599627370496.0;
                double f = __fma_rd(dy1, dw0s, 4503599627370496.0);
                f = __fma_rd(dy0, dw1s, f);
                const u64 bt = (u64)__double_as_longlong(f) & 0xFFFFFFFFFFFFFull;
                const u64 qh = (u64)y1 * w1p + bt;
                x[i] = x[i] * w - qh * q;
            } else if (KIND == 13) {  // one IMAD.WIDE and one DFMA per slot, independent chains: do the two pipes overlap?
                x[i] = (u64)(u32)x[i] * b32 + x[i];
                d[i] = fma(d[i], dw, dq);
            } else if (KIND == 14) {  // hybrid butterfly, values in [0, 8q), one conditional subtraction of 4q
                if (i & 1) {
                    u64 X = x[i - 1], Y = x[i];
                    X = X >= q4 ? X - q4 : X;
                    const u32 y0 = (u32)Y, y1 = (u32)(Y >> 32);
                    const double dy1 = __hiloint2double(0x43300000, (int)y1) - 4503599627370496.0;
                    const double dy0 = __hiloint2double(0x43300000, (int)y0) - 4503599627370496.0;
                    double f = __fma_rd(dy1, dw0s, 4503599627370496.0);
                    f = __fma_rd(dy0, dw1s, f);
                    const u64 bt = (u64)__double_as_longlong(f) & 0xFFFFFFFFFFFFFull;
                    const u64 qh = (u64)y1 * w1p + bt;
                    const u64 T = Y * w - qh * q;
                    x[i - 1] = X + T;
                    x[i] = X - T + q4;
                }
            } else if (KIND == 15) {  // two IMAD.WIDE per DFMA
                x[i] = (u64)(u32)x[i] * b32 + x[i];
                x[i] = (u64)(u32)x[i] * c32 + x[i];
                d[i] = fma(d[i], dw, dq);
            } else if (KIND == 8) {  // uncorrected lazy butterfly
                if (i & 1) {
                    u64 X = x[i - 1], Y = x[i];
                    u64 h = __umul64hi(Y, ws);
                    u64 T = Y * w - h * q;
                    x[i - 1] = X + T;
                    x[i] = X - T + q2;
                }
            }
        }
    }
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) acc += x[i] + a32[i] + (u64)__double_as_longlong(d[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int KIND>
static void run(const char* name, double ops_per_iter_per_thread, u64* d, int sms, double mhz) {
    const int threads = 1024, blocks = sms * 2;
    k<KIND><<<blocks, threads>>>(d, 12345, 0x3ffffffffc0001ull);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k<KIND><<<blocks, threads>>>(d, 12345 + r, 0x3ffffffffc0001ull);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double total = (double)blocks * threads * ITERS * ops_per_iter_per_thread;
    const double per_s = total / (best * 1e-3);
    printf("%-28s %8.3f ms  %8.2f Gop/s  %7.2f lanes/clk/SM (at %.0f MHz nominal)\n", name, best, per_s / 1e9,
           per_s / (sms * mhz * 1e6), mhz);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1e3;
    printf("%s  SMs=%d  clock=%.0f MHz\n", p.name, p.multiProcessorCount, mhz);
    u64* d;
    cudaMalloc(&d, sizeof(u64) * 1024 * p.multiProcessorCount * 2);
    run<0>("IMAD (32-bit mad.lo)", ILP, d, p.multiProcessorCount, mhz);
    run<1>("IMAD.WIDE.U32", ILP, d, p.multiProcessorCount, mhz);
    run<2>("IMAD.HI.U32", ILP, d, p.multiProcessorCount, mhz);
    run<3>("IADD3+LOP3 pair", 2 * ILP, d, p.multiProcessorCount, mhz);
    run<4>("mul.lo.u64 (+add)", ILP, d, p.multiProcessorCount, mhz);
    run<5>("mul.hi.u64 (+add)", ILP, d, p.multiProcessorCount, mhz);
    run<6>("Shoup lazy mulmod", ILP, d, p.multiProcessorCount, mhz);
    run<7>("Harvey butterfly (corrected)", ILP / 2, d, p.multiProcessorCount, mhz);
    run<8>("lazy butterfly (uncorrected)", ILP / 2, d, p.multiProcessorCount, mhz);
    run<9>("DFMA", ILP, d, p.multiProcessorCount, mhz);
    run<10>("FP64 lazy butterfly", ILP / 2, d, p.multiProcessorCount, mhz);
    run<11>("FP64 + integer butterfly pair", ILP / 2, d, p.multiProcessorCount, mhz);
    run<12>("hybrid Shoup mulmod (FP64 cross terms)", ILP, d, p.multiProcessorCount, mhz);
    run<14>("hybrid butterfly [0,8q)", ILP / 2, d, p.multiProcessorCount, mhz);
    run<13>("IMAD.WIDE + DFMA, independent", ILP, d, p.multiProcessorCount, mhz);
    run<15>("2 IMAD.WIDE + DFMA, independent", ILP, d, p.multiProcessorCount, mhz);
    return 0;
}
