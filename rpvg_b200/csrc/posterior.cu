// Log-likelihoods of path-group sets (haplotype combinations) — the `log`-bound core of the haplotype models.
//
// Reference semantics restated (never its code): for a set s of k paths,
//   LL_s = sum_r count_r * log(noise_r + sum_{p in s} P[r, p] / k)
// (rpvg:src/path_estimator.cpp:354-361 Full, :424-440 Bounded, :527-545 Gibbs). The per-set terms that do not touch the
// matrix (log frequencies, log numPermutations, the log-sum-exp normalisation, Bounded's sequential pruning) stay on the
// host side of the ABI (rpvg_b200/host), in the reference's own order, because they are O(#sets), not O(#sets x M).
//
// Layout: problems (clusters) concatenated; per problem a dense column-major matrix rows x paths, as the reference holds
// it, plus noise[rows] and counts[rows]. One warp per set: lanes stride the rows (coalesced column reads, the columns of a
// problem are shared by all of its sets and stay in L2/L1), fp64 log, fixed summation tree (deterministic).
// Bound: fp64 ALU / log throughput, not HBM.

#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "common.cuh"
#include "grouped.cuh"

namespace {

constexpr int kLoglikThreads = 256;
constexpr uint32_t kMaxGroupSizeLoglik = 8;

struct LoglikArgs {

    uint32_t n_problems;
    uint32_t group_size;
    uint64_t n_sets;
    const uint64_t * row_offset;
    const uint32_t * rows;          // rows of each problem when they do not fill [row_offset[p], row_offset[p + 1]); else NULL
    const uint64_t * path_offset;
    const uint64_t * matrix_offset;
    const uint64_t * set_offset;
    const double * matrix;
    const double * noise;
    const double * counts;
    const uint32_t * sets;
    double * loglik;
};

__global__ void __launch_bounds__(kLoglikThreads) groupLoglikKernel(const LoglikArgs args) {

    const uint32_t lane = threadIdx.x & 31;
    const uint64_t set = (blockIdx.x * static_cast<uint64_t>(kLoglikThreads) + threadIdx.x) >> 5;
    if (set >= args.n_sets) { return; }

    const uint32_t p = clusterOf(args.set_offset, args.n_problems, set);
    const uint64_t r0 = args.row_offset[p];
    const uint32_t rows = args.rows ? args.rows[p] : static_cast<uint32_t>(args.row_offset[p + 1] - r0);
    const double * matrix = args.matrix + args.matrix_offset[p];
    const double * noise = args.noise + r0;
    const double * counts = args.counts + r0;

    const uint32_t k = args.group_size;
    const double group_size = static_cast<double>(k);

    // a set member outside the problem's columns cannot be evaluated: the set's result is NaN (nothing is read out of range)
    const uint32_t n_cols = static_cast<uint32_t>(args.path_offset[p + 1] - args.path_offset[p]);
    bool in_range = true;

    const double * cols[kMaxGroupSizeLoglik];
    #pragma unroll
    for (uint32_t q = 0; q < kMaxGroupSizeLoglik; ++q) {
        uint32_t member = (q < k) ? args.sets[set * k + q] : 0;
        if (member >= n_cols) {
            in_range = false;
            member = 0;
        }
        cols[q] = matrix + static_cast<uint64_t>(member) * rows;
    }

    double acc = 0;

    for (uint32_t r = lane; r < rows; r += 32) {

        // noise + col_1 / k + col_2 / k + ... in the reference's order
        double prob = noise[r];
        #pragma unroll
        for (uint32_t q = 0; q < kMaxGroupSizeLoglik; ++q) {
            if (q < k) { prob += cols[q][r] / group_size; }
        }
        acc += counts[r] * log(prob);
    }

    acc = warpSum(acc);
    if (lane == 0) { args.loglik[set] = in_range ? acc : nan(""); }
}

}

extern "C" int rpvg_b200_group_loglik(rpvg_b200_ctx * ctx, const rpvg_b200_loglik_problems * pr, double * loglik) {

    if (!ctx || !pr || !loglik) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (pr->n_problems == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no problems"); }
    if (pr->group_size == 0 || pr->group_size > kMaxGroupSizeLoglik) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "group_size must be 1..8"); }

    const uint32_t P = pr->n_problems;
    const uint64_t R = pr->row_offset[P];
    const uint64_t Mx = pr->matrix_offset[P];
    const uint64_t S = pr->set_offset[P];

    if (S == 0) { return RPVG_B200_OK; }

    for (uint32_t p = 0; p < P; ++p) {
        const uint64_t rows = pr->row_offset[p + 1] - pr->row_offset[p];
        const uint64_t paths = pr->path_offset[p + 1] - pr->path_offset[p];
        if (pr->matrix_offset[p + 1] - pr->matrix_offset[p] != rows * paths) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "matrix_offset does not match rows x paths"); }
        if (rows >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "too many rows in one problem"); }
    }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 4 * pad(8 * (P + 1)) + pad(8 * Mx) + 2 * pad(8 * R) + pad(4 * S * pr->group_size) + pad(8 * S) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_row_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_matrix_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_set_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    double * d_matrix = reinterpret_cast<double *>(take(8 * Mx));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    double * d_counts = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_sets = reinterpret_cast<uint32_t *>(take(4 * S * pr->group_size));
    double * d_loglik = reinterpret_cast<double *>(take(8 * S));

    int status = RPVG_B200_OK;
    cudaError_t err = cudaSuccess;

    #define RPVG_H2D(dst, src, n)                                                                                  \
        if (err == cudaSuccess && (n) > 0) { err = cudaMemcpyAsync((dst), (src), (n), cudaMemcpyHostToDevice, s); }

    RPVG_H2D(d_row_offset, pr->row_offset, 8 * (P + 1))
    RPVG_H2D(d_path_offset, pr->path_offset, 8 * (P + 1))
    RPVG_H2D(d_matrix_offset, pr->matrix_offset, 8 * (P + 1))
    RPVG_H2D(d_set_offset, pr->set_offset, 8 * (P + 1))
    RPVG_H2D(d_matrix, pr->matrix, 8 * Mx)
    RPVG_H2D(d_noise, pr->noise, 8 * R)
    RPVG_H2D(d_counts, pr->counts, 8 * R)
    RPVG_H2D(d_sets, pr->sets, 4 * S * pr->group_size)
    #undef RPVG_H2D

    if (err == cudaSuccess) {

        LoglikArgs args;
        args.n_problems = P;
        args.group_size = pr->group_size;
        args.n_sets = S;
        args.row_offset = d_row_offset;
        args.rows = nullptr;
        args.path_offset = d_path_offset;
        args.matrix_offset = d_matrix_offset;
        args.set_offset = d_set_offset;
        args.matrix = d_matrix;
        args.noise = d_noise;
        args.counts = d_counts;
        args.sets = d_sets;
        args.loglik = d_loglik;

        const uint64_t blocks = (S * 32 + kLoglikThreads - 1) / kLoglikThreads;
        groupLoglikKernel<<<static_cast<unsigned>(blocks), kLoglikThreads, 0, s>>>(args);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess) { err = cudaMemcpyAsync(loglik, d_loglik, 8 * S, cudaMemcpyDeviceToHost, s); }
    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

    if (err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_group_loglik: ") + cudaGetErrorString(err)); }

    cudaFreeAsync(base, s);
    return status;
}

// The same sums on matrices that are already on the device (rpvg_b200_grouped_create): only the sets go up and the
// log-likelihoods come down.
extern "C" int rpvg_b200_grouped_loglik(rpvg_b200_ctx * ctx, const rpvg_b200_grouped * grouped, uint32_t group_size, const uint64_t * set_offset, const uint32_t * sets,
                                        double * loglik) {

    if (!ctx || !grouped || !set_offset || !loglik) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (group_size == 0 || group_size > kMaxGroupSizeLoglik) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "group_size must be 1..8"); }

    const uint32_t P = grouped->n_clusters;
    const uint64_t S = set_offset[P];

    if (S == 0) { return RPVG_B200_OK; }
    if (!sets) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = pad(8 * (P + 1)) + pad(4 * S * group_size) + pad(8 * S) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    uint64_t * d_set_offset = reinterpret_cast<uint64_t *>(base);
    uint32_t * d_sets = reinterpret_cast<uint32_t *>(base + pad(8 * (P + 1)));
    double * d_loglik = reinterpret_cast<double *>(base + pad(8 * (P + 1)) + pad(4 * S * group_size));

    cudaError_t err = cudaMemcpyAsync(d_set_offset, set_offset, 8 * (P + 1), cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) { err = cudaMemcpyAsync(d_sets, sets, 4 * S * group_size, cudaMemcpyHostToDevice, s); }

    if (err == cudaSuccess) {

        LoglikArgs args;
        args.n_problems = P;
        args.group_size = group_size;
        args.n_sets = S;
        args.row_offset = grouped->d_row_offset;
        args.rows = grouped->d_n_classes;
        args.path_offset = grouped->d_group_offset;
        args.matrix_offset = grouped->d_matrix_offset;
        args.set_offset = d_set_offset;
        args.matrix = grouped->d_matrix;
        args.noise = grouped->d_noise;
        args.counts = grouped->d_counts;
        args.sets = d_sets;
        args.loglik = d_loglik;

        const uint64_t blocks = (S * 32 + kLoglikThreads - 1) / kLoglikThreads;
        groupLoglikKernel<<<static_cast<unsigned>(blocks), kLoglikThreads, 0, s>>>(args);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess) { err = cudaMemcpyAsync(loglik, d_loglik, 8 * S, cudaMemcpyDeviceToHost, s); }
    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

    int status = RPVG_B200_OK;
    if (err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_grouped_loglik: ") + cudaGetErrorString(err)); }

    cudaFreeAsync(base, s);
    return status;
}

// Host matrices put on the device once, for several rounds of rpvg_b200_grouped_loglik (problems->sets is not used).
extern "C" int rpvg_b200_grouped_upload(rpvg_b200_ctx * ctx, const rpvg_b200_loglik_problems * pr, rpvg_b200_grouped ** out) {

    if (!ctx || !pr || !out) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    *out = nullptr;
    if (pr->n_problems == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no problems"); }

    const uint32_t P = pr->n_problems;
    const uint64_t R = pr->row_offset[P];
    const uint64_t Mx = pr->matrix_offset[P];

    std::vector<uint32_t> rows(P);

    for (uint32_t p = 0; p < P; ++p) {
        const uint64_t r = pr->row_offset[p + 1] - pr->row_offset[p];
        const uint64_t paths = pr->path_offset[p + 1] - pr->path_offset[p];
        if (pr->matrix_offset[p + 1] - pr->matrix_offset[p] != r * paths) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "matrix_offset does not match rows x paths"); }
        if (r >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "too many rows in one problem"); }
        rows[p] = static_cast<uint32_t>(r);
    }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 3 * pad(8 * (P + 1)) + pad(4 * P) + pad(8 * Mx) + 2 * pad(8 * R) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_row_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_matrix_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint32_t * d_rows = reinterpret_cast<uint32_t *>(take(4 * P));
    double * d_matrix = reinterpret_cast<double *>(take(8 * Mx));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    double * d_counts = reinterpret_cast<double *>(take(8 * R));

    cudaError_t err = cudaSuccess;
    auto up = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); }
    };

    up(d_row_offset, pr->row_offset, 8 * (P + 1));
    up(d_path_offset, pr->path_offset, 8 * (P + 1));
    up(d_matrix_offset, pr->matrix_offset, 8 * (P + 1));
    up(d_rows, rows.data(), 4 * P);
    up(d_matrix, pr->matrix, 8 * Mx);
    up(d_noise, pr->noise, 8 * R);
    up(d_counts, pr->counts, 8 * R);

    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }   // `rows` and the caller's arrays may go away

    if (err != cudaSuccess) {
        cudaFreeAsync(base, s);
        return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_grouped_upload: ") + cudaGetErrorString(err));
    }

    auto * grouped = new rpvg_b200_grouped();
    grouped->n_clusters = P;
    grouped->base = base;
    grouped->d_matrix = d_matrix;
    grouped->d_noise = d_noise;
    grouped->d_counts = d_counts;
    grouped->d_row_offset = d_row_offset;
    grouped->d_group_offset = d_path_offset;
    grouped->d_matrix_offset = d_matrix_offset;
    grouped->d_n_classes = d_rows;
    grouped->h_rows_total = R;
    *out = grouped;

    return RPVG_B200_OK;
}


// ---- fp64 log throughput of this GPU: the roofline denominator of the log-likelihood kernels ---------------------------
namespace {

// Every thread evaluates 4 independent chains of log(): the instruction mix of the pair kernel's inner loop (one add,
// one log, one fused multiply-add per evaluation) without any memory traffic.
__global__ void logPeakKernel(const uint32_t iterations, const double seed, double * __restrict__ sink) {

    double x[4];
    #pragma unroll
    for (int q = 0; q < 4; ++q) { x[q] = seed + 1e-3 * (threadIdx.x * 4 + q) + 1e-6 * blockIdx.x; }

    double acc[4] = {0, 0, 0, 0};

    for (uint32_t it = 0; it < iterations; ++it) {
        #pragma unroll
        for (int q = 0; q < 4; ++q) {
            acc[q] = fma(1.000001, log(x[q] + acc[q] * 1e-9), acc[q]);
        }
    }

    const double total = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    if (total == 12345.678) { sink[0] = total; }   // never true: keeps the loop alive
}

}

namespace {

// Every thread runs 8 independent chains of fused multiply-adds: the fp64 pipe alone.
__global__ void fmaPeakKernel(const uint32_t iterations, const double seed, double * __restrict__ sink) {

    double x[8];
    #pragma unroll
    for (int q = 0; q < 8; ++q) { x[q] = seed + 1e-3 * (threadIdx.x * 8 + q) + 1e-6 * blockIdx.x; }

    const double a = 1.0 - 1e-9, b = 1e-9;

    for (uint32_t it = 0; it < iterations; ++it) {
        #pragma unroll
        for (int q = 0; q < 8; ++q) { x[q] = fma(x[q], a, b); }
    }

    const double total = ((x[0] + x[1]) + (x[2] + x[3])) + ((x[4] + x[5]) + (x[6] + x[7]));
    if (total == 12345.678) { sink[0] = total; }   // never true: keeps the loop alive
}

// best of three timed launches (after one warm-up) of a kernel<<<grid, 256>>>(iterations, seed, sink); returns
// per_thread * iterations * threads / seconds
template <typename Kernel>
int peakOf(rpvg_b200_ctx * ctx, Kernel kernel, const uint32_t iterations, const double per_thread, const char * name, double * per_second) {

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    double * d_sink = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&d_sink), 256, ctx->pool, s));

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaError_t err = cudaEventCreate(&ev0);
    if (err == cudaSuccess) { err = cudaEventCreate(&ev1); }

    const unsigned grid = 8u * static_cast<unsigned>(ctx->num_sms), block = 256;
    double best = 0;

    for (int rep = 0; rep < 4 && err == cudaSuccess; ++rep) {   // first repetition = warm-up

        cudaEventRecord(ev0, s);
        kernel<<<grid, block, 0, s>>>(iterations, 0.5, d_sink);
        ctx->launches++;
        cudaEventRecord(ev1, s);
        err = cudaEventSynchronize(ev1);

        float ms = 0;
        if (err == cudaSuccess) { err = cudaEventElapsedTime(&ms, ev0, ev1); }
        if (err == cudaSuccess && rep > 0 && ms > 0) { best = std::max(best, per_thread * iterations * grid * block / (ms * 1e-3)); }
    }

    if (ev0) { cudaEventDestroy(ev0); }
    if (ev1) { cudaEventDestroy(ev1); }
    cudaFreeAsync(d_sink, s);

    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string(name) + ": " + cudaGetErrorString(err)); }

    *per_second = best;
    return RPVG_B200_OK;
}

}

extern "C" int rpvg_b200_fp64_peak(rpvg_b200_ctx * ctx, double * fma_per_second) {

    if (!ctx || !fma_per_second) { return RPVG_B200_ERR_INVALID; }
    return peakOf(ctx, fmaPeakKernel, 20000, 8.0, "rpvg_b200_fp64_peak", fma_per_second);
}

extern "C" int rpvg_b200_log_peak(rpvg_b200_ctx * ctx, double * logs_per_second) {

    if (!ctx || !logs_per_second) { return RPVG_B200_ERR_INVALID; }
    return peakOf(ctx, logPeakKernel, 2000, 4.0, "rpvg_b200_log_peak", logs_per_second);
}
