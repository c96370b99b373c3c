// emDenseKernel — EM of the small path clusters (sm_100a): one CTA per cluster, the matrix dense in shared memory.
//
// Most path clusters of a run are tiny (a handful of paths and read classes) and many of them need thousands of EM
// iterations. On the sliced-ELL kernels (em_sell.cuh) an iteration of such a cluster costs ~2000 cycles of fixed latency
// (chunk descriptors, lane reductions, a block-wide noise sum, two barriers) whatever its size, and the per-CTA timeline
// (tools/em_timeline.py) showed exactly these clusters as the tail of a batch. Here thread t owns read class t in the
// E-step and path t in the M-step; the noise probability is column N of the same matrix, so there is no reduction at
// all in the loop:
//
//   P[r][j], r < M, j <= N (row stride DIM + 1 doubles: odd, so both the row-wise and the column-wise walk are free of
//   bank conflicts), a[j], w[r] in shared memory.
//   E: w_r = count_r / sum_{j <= N} P[r][j] a[j]             (rpvg:src/path_abundance_estimator.cpp:66-80)
//   M: a_j = a_j * (sum_r P[r][j] w_r) / total, j <= N        (rpvg:src/path_abundance_estimator.cpp:82-98)
//   one barrier after each step; the second one carries the convergence vote.
//
// Zero entries add exactly nothing, so with one running sum (sides 16 and 32) the E-step sum has the bits of the sparse
// sequential one. Deterministic like the other kernels (fixed order, no atomics).
#pragma once

#include "common.cuh"
#include "em_kernels.cuh"
#include "em_sell.cuh"

template <int DIM> struct DenseShape {
    static constexpr int kThreads = DIM < 32 ? 32 : DIM;
    static constexpr int kStride = DIM + 1;
    // running sums per thread: one for the single-warp kernels (measured: the four-way loop's remainder handling costs a
    // short sum more than the 9-cycle dependent chain it saves: 0.74 vs 0.45 us per iteration at side 16)
    static constexpr int kWays = DIM > 32 ? 4 : 1;
    static constexpr size_t kSmemBytes = sizeof(double) * (static_cast<size_t>(DIM) * kStride + 2 * DIM);
};

// sum_k x[k * x_stride] * y[k], k < n
template <int WAYS>
__device__ __forceinline__ double denseDot(const double * x, const uint32_t x_stride, const double * y, const uint32_t n) {

    if (WAYS == 1) {

        double s = 0;
        #pragma unroll 4
        for (uint32_t k = 0; k < n; ++k) { s = fma(x[k * x_stride], y[k], s); }
        return s;
    }

    double s[WAYS];
    #pragma unroll
    for (int u = 0; u < WAYS; ++u) { s[u] = 0; }

    uint32_t k = 0;
    for (; k + WAYS <= n; k += WAYS) {
        #pragma unroll
        for (int u = 0; u < WAYS; ++u) { s[u] = fma(x[(k + u) * x_stride], y[k + u], s[u]); }
    }
    for (; k < n; ++k) { s[0] = fma(x[k * x_stride], y[k], s[0]); }

    double total = s[0];
    #pragma unroll
    for (int u = 1; u < WAYS; ++u) { total += s[u]; }
    return total;
}

// The two barriers of an iteration; a CTA of one warp needs no hardware barrier for them.
template <int THREADS>
__device__ __forceinline__ void denseBarrier() {

    if (THREADS == 32) { __syncwarp(); } else { __syncthreads(); }
}

template <int THREADS>
__device__ __forceinline__ bool denseVote(const bool ok) {

    if (THREADS == 32) {
        const bool all = __all_sync(0xffffffffu, ok);
        __syncwarp();
        return all;
    }
    return __syncthreads_and(ok) != 0;
}

template <int DIM>
__global__ void __launch_bounds__(DenseShape<DIM>::kThreads) emDenseKernel(const EmArgs args, const uint32_t * __restrict__ clusters) {

    typedef DenseShape<DIM> Shape;
    constexpr int THREADS = Shape::kThreads;
    constexpr uint32_t stride = Shape::kStride;

    extern __shared__ __align__(16) unsigned char dense_smem[];
    __shared__ double red[32];

    double * P = reinterpret_cast<double *>(dense_smem);   // [DIM][DIM + 1]
    double * a = P + DIM * stride;                         // [DIM]
    double * w = a + DIM;                                  // [DIM]

    const uint32_t tid = threadIdx.x;

    if (*args.error) { return; }

#ifdef RPVG_B200_EM_TIMELINE
    unsigned long long timeline_t0, timeline_t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t0));
#endif

    const uint32_t c = clusters[blockIdx.x];

    const uint64_t r0 = args.read_offset[c];
    const uint32_t M = static_cast<uint32_t>(args.read_offset[c + 1] - r0);
    const uint64_t p0 = args.path_offset[c];
    const uint32_t N = static_cast<uint32_t>(args.path_offset[c + 1] - p0);

    if (M == 0) {   // rpvg:src/path_abundance_estimator.cpp:20-22 — only the reset runs

        for (uint32_t j = tid; j < N; j += THREADS) { args.abundances[p0 + j] = 0; }

        if (tid == 0) {
            args.noise_count[c] = 0;
            args.total_count[c] = 0;
            args.iterations[c] = 0;
        }
        return;
    }

    for (uint32_t i = tid; i < M * stride; i += THREADS) { P[i] = 0; }
    __syncthreads();

    double count = 0;

    if (tid < M) {

        const uint64_t * row_ptr = args.row_ptr + r0 + tid;
        const uint64_t e1 = row_ptr[1];
        double * row = P + tid * stride;

        for (uint64_t e = row_ptr[0]; e < e1; ++e) { row[args.col[e]] = args.val[e]; }
        row[N] = args.noise[r0 + tid];

        count = args.cnt[r0 + tid];
    }

    // rpvg:src/path_abundance_estimator.cpp:54 — single-precision quotient widened to double. (A path without entries
    // drops to 0 in the first update and stays out of the convergence test, like a path that starts at 0.)
    double a_own = static_cast<double>(1 / static_cast<float>(N + 1));
    if (tid <= N) { a[tid] = a_own; }

    const double total = blockSum(count, red);   // (has the barriers the fill needs)
    const double inv_total = 1.0 / total;

    const double conv = args.params.max_rel_em_conv;
    const double min_abundance = args.params.min_em_abundance;

    uint32_t conv_its = 0;
    uint32_t it = 0;

#ifdef RPVG_B200_EM_TIMELINE
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t1));
#endif

    while (it < args.params.max_em_its) {

        if (tid < M) { w[tid] = count * sellReciprocal(denseDot<Shape::kWays>(P + tid * stride, 1, a, N + 1)); }

        denseBarrier<THREADS>();

        bool ok = true;

        if (tid <= N) {

            const double g = denseDot<Shape::kWays>(P + tid, stride, w, M);
            const double a_new = (a_own * g) * inv_total;

            if (a_new >= min_abundance && fabs(a_new - a_own) > conv * a_new) { ok = false; }

            a[tid] = a_new;
            a_own = a_new;
        }

        ++it;

        if (denseVote<THREADS>(ok)) {
            if (++conv_its == args.params.min_em_conv_its) { break; }
        } else {
            conv_its = 0;
        }
    }

    // rpvg:src/path_abundance_estimator.cpp:100-113
    double noise_part = 0;

    if (tid < N) {
        if (a_own < min_abundance) {
            noise_part = a_own * total;
            args.abundances[p0 + tid] = 0;
        } else {
            args.abundances[p0 + tid] = a_own * total;
        }
    }

    const double low_sum = blockSum(noise_part, red);

    if (tid == 0) {
        args.noise_count[c] = low_sum + a[N] * total;
        args.total_count[c] = total;
        args.iterations[c] = it;
    }

#ifdef RPVG_B200_EM_TIMELINE
    if (tid == 0) {
        unsigned long long timeline_t2;
        uint32_t smid;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t2));
        asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        const unsigned int slot = atomicAdd(&rpvg_timeline_count, 1u);
        if (slot < 65536) {
            unsigned long long * rec = rpvg_timeline_records + 12ull * slot;
            rec[0] = DIM; rec[1] = 1; rec[2] = 0; rec[3] = c; rec[4] = smid; rec[5] = M; rec[6] = N;
            rec[7] = timeline_t0; rec[8] = timeline_t1; rec[9] = timeline_t2; rec[10] = it; rec[11] = 1;
        }
    }
#endif
}
