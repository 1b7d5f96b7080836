// How many clusters of a given shape are co-resident on this GPU (cudaOccupancyMaxActiveClusters): the NTT cluster kernel's
// shapes at N = 2^14 (8 CTAs x 128 threads x 16 KiB) and N = 2^16 (16 x 256 x 32 KiB, or 8 x 256 x 64 KiB).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/cluster_occ tools/cluster_occ.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int REGS_HINT>
__global__ void __launch_bounds__(256, REGS_HINT) k(float* p) {
    extern __shared__ float sm[];
    sm[threadIdx.x] = p[threadIdx.x];
    __syncthreads();
    p[threadIdx.x] = sm[(threadIdx.x + 1) % blockDim.x];
}
template <class K>
static void probe(K kern, const char* name, int cl, int threads, int smem) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (cl > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cl * 1024);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cl;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem);
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    printf("%-34s cluster %2d x %3d thr x %3d KiB: %4d clusters = %5d CTAs of %5d slots (%d/SM x %d SMs) = %.1f %%  %s\n", name, cl, threads,
           smem >> 10, n, n * cl, per_sm * pr.multiProcessorCount, per_sm, pr.multiProcessorCount,
           100.0 * n * cl / (per_sm * pr.multiProcessorCount), e == cudaSuccess ? "" : cudaGetErrorString(e));
}
int main() {
    probe(k<4>, "N=2^16 default (16 CTAs)", 16, 256, 32 << 10);
    probe(k<3>, "N=2^16 alternative (8 CTAs)", 8, 256, 64 << 10);
    probe(k<4>, "N=2^15 (8 CTAs)", 8, 256, 32 << 10);
    probe(k<8>, "N=2^14 (8 CTAs x 128)", 8, 128, 16 << 10);
    probe(k<8>, "N=2^14 (4 CTAs x 256)", 4, 256, 32 << 10);
    probe(k<4>, "cluster of 4 x 256 x 32 KiB", 4, 256, 32 << 10);
    probe(k<4>, "cluster of 2 x 256 x 32 KiB", 2, 256, 32 << 10);
    return 0;
}
