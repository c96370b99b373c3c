// Shared declarations of librpvg_b200.so (sm_100a only).
#pragma once

#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "rpvg_b200.h"

constexpr int kNumSideStreams = 12;

struct rpvg_b200_ctx {

    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int num_sms = 148;
    int max_smem_optin = 0;
    uint64_t launches = 0;
    std::string error;

    // Every device allocation of the library comes out of this pool (cudaMallocFromPoolAsync) and goes back to it with
    // cudaFreeAsync: freed batch memory stays with the ctx for the next batch, and the device's default pool — which
    // belongs to the host application — is left alone.
    cudaMemPool_t pool = nullptr;

    // EM launch machinery (em.cu): side streams for concurrent bins, largest thread-block cluster this device accepts
    std::vector<cudaStream_t> side_streams;   // kNumSideStreams: pairs of equal priority, the last four lowest
    std::vector<cudaEvent_t> join_events;
    std::vector<cudaStream_t> lane_streams;   // highest priority; the largest groups run here one cluster per launch
    std::vector<cudaEvent_t> lane_events;
    cudaEvent_t fork_event = nullptr;
    uint32_t max_group_size = 1;
    uint32_t stream_grid = 0;                 // CTAs of a cooperative emStreamKernel launch that are resident at once

    // rpvg_b200_pinned_scratch: grow-only page-locked host buffers, one per slot
    static constexpr uint32_t kPinnedSlots = 8;
    void * pinned[kPinnedSlots] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    uint64_t pinned_bytes[kPinnedSlots] = {0, 0, 0, 0, 0, 0, 0, 0};
};

// em.cu: host-side shape checks of a cluster arena (offsets start at 0 and never decrease — over all read classes, not
// only at cluster boundaries —, no cluster without paths, sizes within the 32-bit limits of the device code).
int rpvgCheckClusters(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters);

// em.cu: sets the kernel attributes of the EM kernels and probes the largest usable thread-block cluster.
int rpvgEmInit(rpvg_b200_ctx * ctx);

inline int rpvgFail(rpvg_b200_ctx * ctx, int status, const std::string & message) {

    if (ctx) { ctx->error = message; }
    return status;
}

#define RPVG_CUDA(ctx, call)                                                                                     \
    do {                                                                                                         \
        cudaError_t rpvg_cuda_err = (call);                                                                      \
        if (rpvg_cuda_err != cudaSuccess) {                                                                      \
            return rpvgFail((ctx), RPVG_B200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(rpvg_cuda_err)); \
        }                                                                                                        \
    } while (0)

// Launch bookkeeping: bench.py reports gpu_launches from this counter.
#define RPVG_LAUNCHED(ctx)                                                                                       \
    do {                                                                                                         \
        (ctx)->launches++;                                                                                       \
        cudaError_t rpvg_launch_err = cudaGetLastError();                                                        \
        if (rpvg_launch_err != cudaSuccess) {                                                                    \
            return rpvgFail((ctx), RPVG_B200_ERR_CUDA, std::string("kernel launch: ") + cudaGetErrorString(rpvg_launch_err)); \
        }                                                                                                        \
    } while (0)

// Returns a stream-ordered allocation on every way out of a function (the RPVG_CUDA / RPVG_LAUNCHED early returns included).
struct DeviceAllocationGuard {

    cudaStream_t stream;
    void * ptr;

    ~DeviceAllocationGuard() {
        if (ptr) { cudaFreeAsync(ptr, stream); }
    }
};

// Device error flags a kernel may raise (checked at download).
enum : uint32_t {
    RPVG_DEV_ERR_COL_RANGE = 1u,   // col[e] >= N_c
    RPVG_DEV_ERR_COL_ORDER = 2u,   // entries of a read class not strictly ascending by path index
    RPVG_DEV_ERR_VALUE = 4u,       // non-finite or negative probability
    RPVG_DEV_ERR_CAPACITY = 8u     // internal: a slice outgrew its scratch area
};

#ifdef __CUDACC__

__device__ __forceinline__ double warpSum(double v) {

    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, off);
    }
    return v;
}

// Deterministic block-wide sum (fixed tree). scratch: >= 32 doubles of shared memory. All threads get the result.
__device__ __forceinline__ double blockSum(double v, double * scratch) {

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int num_warps = (blockDim.x + 31) >> 5;

    v = warpSum(v);
    __syncthreads();   // scratch may still be read from a previous call
    if (lane == 0) { scratch[warp] = v; }
    __syncthreads();

    double r = (lane < num_warps) ? scratch[lane] : 0.0;
    r = warpSum(r);
    return r;
}

// First cluster c with offset[c + 1] > idx (offset non-decreasing, offset[n] > idx).
__device__ __forceinline__ uint32_t clusterOf(const uint64_t * __restrict__ offset, uint32_t n, uint64_t idx) {

    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(offset + mid + 1) > idx) { hi = mid; } else { lo = mid + 1; }
    }
    return lo;
}

#endif
