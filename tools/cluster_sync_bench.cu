// Micro-benchmark (tuning aid): cost of the per-iteration exchange pattern of the EM kernel inside a thread-block
// cluster: (a) two cluster.sync() only, (b) + reduce-scatter/all-gather of n doubles over DSMEM, (c) arrive/wait split.
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

template <int MODE>
__global__ void k(int iters, int n, double * out, long long * cycles) {
    extern __shared__ double sm[];
    double * a = sm;
    double * g = sm + n;
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned S = cluster.num_blocks(), rank = cluster.block_rank();
    for (int j = threadIdx.x; j < n; j += blockDim.x) { a[j] = 1.0 / n; g[j] = 1.0; }
    cluster.sync();
    const int j0 = (n * rank) / S, j1 = (n * (rank + 1)) / S;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) { cluster.sync(); cluster.sync(); }
        if (MODE == 1 || MODE == 2) {
            for (int j = threadIdx.x; j < n; j += blockDim.x) { g[j] = a[j] * 1.0001; }
            if (MODE == 1) { cluster.sync(); } else { asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory"); }
            for (int j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
                double p[16];
                #pragma unroll
                for (unsigned s = 0; s < 16; ++s) { p[s] = s < S ? *(cluster.map_shared_rank(g, s) + j) : 0.0; }
                double sum = 0;
                #pragma unroll
                for (unsigned s = 0; s < 16; ++s) { sum += p[s]; }
                sum /= S;
                #pragma unroll
                for (unsigned s = 0; s < 16; ++s) { if (s < S) { *(cluster.map_shared_rank(a, s) + j) = sum; } }
            }
            if (MODE == 1) { cluster.sync(); } else { asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory"); }
        }
        if (MODE == 3) { __syncthreads(); __syncthreads(); }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && rank == 0) { cycles[blockIdx.x / S] = t1 - t0; out[blockIdx.x / S] = a[0]; }
    cluster.sync();
}

template <int MODE>
void run(int S, int threads, int n, int iters) {
    double * out; long long * cyc;
    cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 1024 * 8);
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(S); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = 2 * n * 8 + 1024;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = S; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaLaunchKernelEx(&cfg, k<MODE>, iters, n, out, cyc);
    cudaEventRecord(e0);
    cudaError_t err = cudaLaunchKernelEx(&cfg, k<MODE>, iters, n, out, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("mode %d S %2d threads %3d n %5d: %8.3f us/iter  %7.0f cycles/iter  (%s)\n", MODE, S, threads, n, 1e3 * ms / iters, (double) h / iters, cudaGetErrorString(err));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    const int iters = 2000;
    for (int S : {1, 2, 4, 8, 16}) for (int th : {128, 512}) run<0>(S, th, 64, iters);
    for (int th : {128, 512}) run<3>(1, th, 64, iters);
    for (int S : {2, 4, 8, 16}) for (int n : {64, 512, 2048, 4800}) run<1>(S, 512, n, iters);
    for (int S : {2, 8, 16}) for (int n : {512, 4800}) run<2>(S, 512, n, iters);
    return 0;
}
