// Dependent-issue latency and per-SM throughput of fp64 FMA/ADD, shared-memory loads and fp64 shuffles on this GPU.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/fp64_latency tools/fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void dfmaKernel(double * out, long long * cycles, double x, int n) {
    double acc[CHAINS];
    for (int c = 0; c < CHAINS; ++c) { acc[c] = threadIdx.x + c; }
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        #pragma unroll
        for (int c = 0; c < CHAINS; ++c) { acc[c] = fma(acc[c], x, 1e-9); }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int c = 0; c < CHAINS; ++c) { s += acc[c]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) { *cycles = t1 - t0; }
}

__global__ void ldsChainKernel(unsigned * out, long long * cycles, int n) {
    __shared__ unsigned next[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) { next[i] = (i * 33 + 7) & 1023; }
    __syncthreads();
    unsigned p = threadIdx.x;
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) { p = next[p]; }
    const long long t1 = clock64();
    out[threadIdx.x] = p;
    if (threadIdx.x == 0) { *cycles = t1 - t0; }
}

__global__ void shflChainKernel(double * out, long long * cycles, int n) {
    double v = threadIdx.x;
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) { v += __shfl_xor_sync(0xffffffffu, v, 1); }
    const long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) { *cycles = t1 - t0; }
}

__global__ void rcpChainKernel(double * out, long long * cycles, int n) {
    double v = threadIdx.x + 1.5;
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        double r;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(v));
        double e = fma(-v, r, 1.0);
        r = fma(r, e, r);
        e = fma(-v, r, 1.0);
        r = fma(r, e, r);
        v = r + 1.25;
    }
    const long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) { *cycles = t1 - t0; }
}

int main() {
    double * out; long long * cyc; long long h;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cyc, 8);
    const int n = 4096;
    int threads[] = {32, 128, 512, 1024};
    for (int t : threads) {
        dfmaKernel<1><<<1, t>>>(out, cyc, 0.999, n); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DFMA 1 chain,  %4d threads: %.1f cycles per dependent op\n", t, double(h) / n);
        dfmaKernel<4><<<1, t>>>(out, cyc, 0.999, n); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DFMA 4 chains, %4d threads: %.1f cycles per round of 4 (%.2f warp-DFMA/cycle/SM)\n", t, double(h) / n, 4.0 * (t / 32) * n / double(h));
        dfmaKernel<8><<<1, t>>>(out, cyc, 0.999, n); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DFMA 8 chains, %4d threads: %.1f cycles per round of 8 (%.2f warp-DFMA/cycle/SM)\n", t, double(h) / n, 8.0 * (t / 32) * n / double(h));
    }
    ldsChainKernel<<<1, 32>>>((unsigned *) out, cyc, n); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("LDS dependent chain: %.1f cycles\n", double(h) / n);
    shflChainKernel<<<1, 32>>>(out, cyc, n); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("fp64 shfl_xor + DADD chain: %.1f cycles per level\n", double(h) / n);
    rcpChainKernel<<<1, 32>>>(out, cyc, n); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("reciprocal (rcp.approx + 2 Newton) + DADD chain: %.1f cycles\n", double(h) / n);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
