// EM kernels (sm_100a). See em.cu for the data layout and the reference semantics they restate.
//
// emSellBuildKernel / emSellRunKernel (em_sell.cuh): the EM kernels proper.
// emClusterKernel (here): fallback, one CTA per path cluster with every array in global memory (any N).
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

struct EmArgs {

    const uint32_t * order;
    const uint64_t * read_offset;
    const uint64_t * path_offset;
    const uint64_t * row_ptr;
    const uint32_t * col;
    const double * val;
    const double * noise;
    const double * cnt;
    const uint64_t * col_ptr;
    const uint32_t * csc_row;
    const double * csc_val;
    double * abund;
    double * w;
    double * abundances;
    double * noise_count;
    double * total_count;
    uint32_t * iterations;
    const uint32_t * error;   // device error flags raised by the build kernels; EM is skipped when set
    rpvg_b200_em_params params;
};

constexpr int kEmThreads = 256;
constexpr uint32_t kMaxGroupSize = 16;

template <typename RowPtr>
__device__ __forceinline__ double emRowsStep(const int tid, const int threads, const uint32_t m, const RowPtr * __restrict__ rptr, const uint64_t ebase,
                                             const uint32_t * __restrict__ col, const double * __restrict__ val, const double * __restrict__ noise,
                                             const double * __restrict__ cnt, const double * a, const double a_noise, double * w) {

    double noise_acc = 0;

    for (uint32_t i = tid; i < m; i += threads) {

        const uint64_t e1 = rptr[i + 1] - ebase;
        double s = 0;
        for (uint64_t e = rptr[i] - ebase; e < e1; ++e) { s += val[e] * a[col[e]]; }
        s += noise[i] * a_noise;

        const double wi = cnt[i] / s;
        w[i] = wi;
        noise_acc += noise[i] * wi;
    }

    return noise_acc;
}

// ---- fallback: one CTA per cluster, all arrays in global memory -----------------------------------------------------
__global__ void __launch_bounds__(kEmThreads) emClusterKernel(const EmArgs args) {

    __shared__ double red[32];

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    constexpr int num_warps = kEmThreads / 32;

    if (*args.error) { return; }

    const uint32_t c = args.order[blockIdx.x];

    const uint64_t r0 = args.read_offset[c];
    const uint32_t M = static_cast<uint32_t>(args.read_offset[c + 1] - r0);
    const uint64_t p0 = args.path_offset[c];
    const uint32_t N = static_cast<uint32_t>(args.path_offset[c + 1] - p0);

    double * abundances = args.abundances + p0;

    if (M == 0) {

        for (uint32_t j = tid; j < N; j += kEmThreads) { abundances[j] = 0; }

        if (tid == 0) {
            args.noise_count[c] = 0;
            args.total_count[c] = 0;
            args.iterations[c] = 0;
        }
        return;
    }

    const uint64_t * row_ptr = args.row_ptr + r0;
    const double * noise = args.noise + r0;
    const double * cnt = args.cnt + r0;
    const uint64_t * col_ptr = args.col_ptr + p0;

    double * a = args.abund + p0 + c;   // N + 1 relative abundances, noise last
    double * w = args.w + r0;

    double count_sum = 0;
    for (uint32_t i = tid; i < M; i += kEmThreads) { count_sum += cnt[i]; }
    const double total = blockSum(count_sum, red);

    const double a_init = static_cast<double>(1 / static_cast<float>(N + 1));
    for (uint32_t j = tid; j <= N; j += kEmThreads) { a[j] = a_init; }
    __syncthreads();

    const double conv = args.params.max_rel_em_conv;
    const double min_abundance = args.params.min_em_abundance;

    uint32_t conv_its = 0;
    uint32_t it = 0;

    while (it < args.params.max_em_its) {

        const double a_noise = a[N];
        const double noise_acc = emRowsStep<uint64_t>(tid, kEmThreads, M, row_ptr, 0, args.col, args.val, noise, cnt, a, a_noise, w);
        const double noise_g = blockSum(noise_acc, red);

        bool ok = true;

        for (uint32_t j = warp; j < N; j += num_warps) {

            const uint64_t k1 = col_ptr[j + 1];
            double gsum = 0;
            for (uint64_t k = col_ptr[j] + lane; k < k1; k += 32) { gsum += args.csc_val[k] * w[args.csc_row[k]]; }
            gsum = warpSum(gsum);

            if (lane == 0) {
                const double a_old = a[j];
                const double a_new = (a_old * gsum) / total;
                a[j] = a_new;
                if (a_new >= min_abundance && fabs(a_new - a_old) / a_new > conv) { ok = false; }
            }
        }

        if (tid == 0) {
            const double a_new = (a_noise * noise_g) / total;
            a[N] = a_new;
            if (a_new >= min_abundance && fabs(a_new - a_noise) / a_new > conv) { ok = false; }
        }

        ++it;

        if (__syncthreads_and(ok)) {
            if (++conv_its == args.params.min_em_conv_its) { break; }
        } else {
            conv_its = 0;
        }
    }

    double noise_part = 0;
    for (uint32_t j = tid; j < N; j += kEmThreads) {

        const double aj = a[j];
        if (aj < min_abundance) {
            noise_part += aj * total;
            abundances[j] = 0;
        } else {
            abundances[j] = aj * total;
        }
    }

    const double low_sum = blockSum(noise_part, red);

    if (tid == 0) {
        args.noise_count[c] = low_sum + a[N] * total;
        args.total_count[c] = total;
        args.iterations[c] = it;
    }
}
