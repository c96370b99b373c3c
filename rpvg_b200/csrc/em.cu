// Matrix build + EM abundance estimation for batches of path clusters (transcripts model and every sub-EM of the
// strains / haplotype-transcripts models).
//
// Reference semantics restated here (never its code): rpvg:src/path_estimator.cpp:55-77 (scatter of read-class
// probabilities into the matrix), :156-166 (row normalisation, (1 - noise) scaling, NaN -> 0, noise as last column) and
// rpvg:src/path_abundance_estimator.cpp:47-114 (EM with the float initialiser, the relative-step convergence rule held
// for min_em_conv_its consecutive iterations, and the final min-abundance cut that moves mass to noise_count).
//
// Device layout (all in one HBM allocation per batch):
//   CSR  : row_ptr[u64], col[u32], val[f64]        one entry per (read class, path) non-zero, normalised
//   CSC  : col_ptr[u64], csc_row[u32], csc_val[f64] the same non-zeros ordered by (global path, read class)
//   rows : noise[f64], cnt[f64]                     the noise column is dense and kept apart
// The reference is dense and column-major; here the matrix is sparse in both orientations so that the E-step (row sums)
// and the M-step (column sums) are both gathers with a fixed summation order: no atomics, results are run-to-run
// deterministic, which matters because the stopping rule is a 1e-3 step test while parity is 1e-6.

#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int kEmThreads = 256;

struct DevicePool {

    // One cudaMallocAsync region, carved in 256-byte aligned pieces.
    char * base = nullptr;
    size_t size = 0;
    size_t used = 0;

    template <typename T>
    T * take(size_t n) {
        used = (used + 255) & ~size_t(255);
        T * p = reinterpret_cast<T *>(base + used);
        used += n * sizeof(T);
        return p;
    }
};

template <typename T>
size_t padded(size_t n) { return ((n * sizeof(T)) + 255) & ~size_t(255); }

}

struct rpvg_b200_batch {

    uint32_t n_clusters = 0;
    uint64_t n_reads = 0;
    uint64_t n_paths = 0;
    uint64_t nnz = 0;

    std::vector<uint64_t> h_read_offset;
    std::vector<uint64_t> h_path_offset;
    std::vector<uint64_t> h_cluster_nnz;   // path non-zeros per cluster
    std::vector<uint32_t> h_iterations;

    DevicePool pool;

    // inputs
    uint64_t * d_read_offset = nullptr;
    uint64_t * d_path_offset = nullptr;
    uint64_t * d_row_ptr = nullptr;
    uint32_t * d_col = nullptr;
    double * d_prob = nullptr;
    double * d_noise = nullptr;
    uint32_t * d_count = nullptr;
    uint32_t * d_order = nullptr;

    // built
    double * d_val = nullptr;
    double * d_cnt = nullptr;
    uint32_t * d_entry_row = nullptr;
    uint32_t * d_key_in = nullptr;
    uint32_t * d_key_out = nullptr;
    uint32_t * d_idx_in = nullptr;
    uint32_t * d_idx_out = nullptr;
    uint64_t * d_col_ptr = nullptr;
    uint32_t * d_csc_row = nullptr;
    double * d_csc_val = nullptr;
    void * d_sort_temp = nullptr;
    size_t sort_temp_bytes = 0;
    int sort_bits = 32;

    // EM state + results
    double * d_abund = nullptr;   // [n_paths + n_clusters]: per cluster N_c + 1 relative abundances (noise last)
    double * d_w = nullptr;       // [n_reads]
    double * d_abundances = nullptr;
    double * d_noise_count = nullptr;
    double * d_total_count = nullptr;
    uint32_t * d_iterations = nullptr;
    uint32_t * d_error = nullptr;

    bool have_results = false;

    // device timestamps around the two phases of rpvg_b200_batch_em (bench.py's roofline numbers)
    cudaEvent_t ev_start = nullptr;
    cudaEvent_t ev_built = nullptr;
    cudaEvent_t ev_done = nullptr;
};

namespace {

// ---- build: one thread per read class -------------------------------------------------------------------------------
// Row normalisation exactly as the reference orders it: v = p / rowsum (left-to-right sum over ascending path index,
// which is what the dense row-wise sum does since zeros add nothing), then v * (1 - noise), NaN -> 0.
__global__ void buildRowsKernel(uint32_t n_clusters, uint64_t n_reads,
                                const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ path_offset,
                                const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col,
                                const double * __restrict__ prob, const double * __restrict__ noise, const uint32_t * __restrict__ count,
                                double * __restrict__ val, double * __restrict__ cnt, uint32_t * __restrict__ entry_row,
                                uint32_t * __restrict__ key, uint32_t * __restrict__ idx, uint32_t * __restrict__ error) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_reads) { return; }

    const uint32_t c = clusterOf(read_offset, n_clusters, r);
    const uint64_t path_base = path_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - path_base;
    const uint32_t local_row = static_cast<uint32_t>(r - read_offset[c]);

    const uint64_t e0 = row_ptr[r];
    const uint64_t e1 = row_ptr[r + 1];

    double row_sum = 0;
    uint32_t err = 0;
    int64_t prev_col = -1;

    for (uint64_t e = e0; e < e1; ++e) {

        const double p = prob[e];
        const uint32_t j = col[e];

        if (j >= num_paths) { err |= RPVG_DEV_ERR_COL_RANGE; }
        if (static_cast<int64_t>(j) <= prev_col) { err |= RPVG_DEV_ERR_COL_ORDER; }
        if (!(p >= 0) || isinf(p)) { err |= RPVG_DEV_ERR_VALUE; }

        prev_col = j;
        row_sum += p;
    }

    const double scale = 1 - noise[r];

    for (uint64_t e = e0; e < e1; ++e) {

        double v = (prob[e] / row_sum) * scale;
        if (isnan(v)) { v = 0; }

        val[e] = v;
        entry_row[e] = local_row;
        key[e] = static_cast<uint32_t>(path_base + min(static_cast<uint64_t>(col[e]), num_paths - 1));
        idx[e] = static_cast<uint32_t>(e);
    }

    cnt[r] = static_cast<double>(count[r]);

    if (err) { atomicOr(error, err); }
}

// col_ptr[j] = first position in the sorted keys with key >= j (j = 0..n_paths).
__global__ void columnPointerKernel(uint64_t n_paths, uint64_t nnz, const uint32_t * __restrict__ sorted_key, uint64_t * __restrict__ col_ptr) {

    const uint64_t j = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (j > n_paths) { return; }

    uint64_t lo = 0, hi = nnz;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (__ldg(sorted_key + mid) >= j) { hi = mid; } else { lo = mid + 1; }
    }
    col_ptr[j] = lo;
}

__global__ void gatherColumnsKernel(uint64_t nnz, const uint32_t * __restrict__ sorted_idx, const uint32_t * __restrict__ entry_row,
                                    const double * __restrict__ val, uint32_t * __restrict__ csc_row, double * __restrict__ csc_val) {

    const uint64_t k = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (k >= nnz) { return; }

    const uint32_t e = sorted_idx[k];
    csc_row[k] = entry_row[e];
    csc_val[k] = val[e];
}

// ---- EM: one CTA per cluster, all iterations inside the kernel ------------------------------------------------------
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
    rpvg_b200_em_params params;
};

__global__ void __launch_bounds__(kEmThreads) emClusterKernel(const EmArgs args) {

    __shared__ double red[32];

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    constexpr int num_warps = kEmThreads / 32;

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

    // rpvg:src/path_abundance_estimator.cpp:54 — single-precision quotient widened to double.
    const double a_init = static_cast<double>(1 / static_cast<float>(N + 1));
    for (uint32_t j = tid; j <= N; j += kEmThreads) { a[j] = a_init; }
    __syncthreads();

    const double conv = args.params.max_rel_em_conv;
    const double min_abundance = args.params.min_em_abundance;

    uint32_t conv_its = 0;
    uint32_t it = 0;

    while (it < args.params.max_em_its) {

        // E-step: w_i = count_i / sum_j p_ij a_j (noise column last, as in the dense left-to-right row sum).
        const double a_noise = a[N];
        double noise_acc = 0;

        for (uint32_t i = tid; i < M; i += kEmThreads) {

            const uint64_t e1 = row_ptr[i + 1];
            double s = 0;
            for (uint64_t e = row_ptr[i]; e < e1; ++e) { s += args.val[e] * a[args.col[e]]; }
            s += noise[i] * a_noise;

            const double wi = cnt[i] / s;
            w[i] = wi;
            noise_acc += noise[i] * wi;
        }

        const double noise_g = blockSum(noise_acc, red);

        // M-step: a_j <- a_j * sum_i p_ij w_i / total, one warp per column.
        bool ok = true;

        for (uint32_t j = warp; j < N; j += num_warps) {

            const uint64_t k1 = col_ptr[j + 1];
            double g = 0;
            for (uint64_t k = col_ptr[j] + lane; k < k1; k += 32) { g += args.csc_val[k] * w[args.csc_row[k]]; }
            g = warpSum(g);

            if (lane == 0) {
                const double a_old = a[j];
                const double a_new = (a_old * g) / total;
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

    // rpvg:src/path_abundance_estimator.cpp:100-113
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

int checkClusters(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl) {

    if (!cl) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "clusters is NULL"); }
    if (cl->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "empty batch (n_clusters == 0)"); }
    if (!cl->read_offset || !cl->path_offset || !cl->row_ptr) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "offset arrays are NULL"); }
    if (cl->read_offset[0] != 0 || cl->path_offset[0] != 0 || cl->row_ptr[0] != 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "offset arrays must start at 0"); }

    for (uint32_t c = 0; c < cl->n_clusters; ++c) {
        if (cl->read_offset[c + 1] < cl->read_offset[c] || cl->path_offset[c + 1] < cl->path_offset[c]) {
            return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "offset arrays must be non-decreasing (cluster " + std::to_string(c) + ")");
        }
        if (cl->path_offset[c + 1] == cl->path_offset[c]) {
            return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "cluster " + std::to_string(c) + " has no paths");
        }
    }

    const uint64_t n_reads = cl->read_offset[cl->n_clusters];
    const uint64_t n_paths = cl->path_offset[cl->n_clusters];
    const uint64_t nnz = cl->row_ptr[n_reads];

    if (n_paths + cl->n_clusters >= 0xffffffffull || nnz >= 0xffffffffull || n_reads >= 0xffffffffull) {
        return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "batch too large: paths, read classes and non-zeros must each stay below 2^32 per batch");
    }
    if (n_reads > 0 && (!cl->noise || !cl->count)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "noise/count arrays are NULL"); }
    if (nnz > 0 && (!cl->col || !cl->prob)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "col/prob arrays are NULL"); }

    return RPVG_B200_OK;
}

}

extern "C" {

void rpvg_b200_em_params_default(rpvg_b200_em_params * params) {

    params->max_em_its = 10000;
    params->max_rel_em_conv = 0.001;
    params->min_em_conv_its = 10;
    params->min_em_abundance = 1e-8;
}

int rpvg_b200_batch_upload(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, rpvg_b200_batch ** out_batch) {

    if (!ctx || !out_batch) { return RPVG_B200_ERR_INVALID; }
    *out_batch = nullptr;

    int status = checkClusters(ctx, cl);
    if (status != RPVG_B200_OK) { return status; }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));

    auto * b = new rpvg_b200_batch();
    b->n_clusters = cl->n_clusters;
    b->n_reads = cl->read_offset[cl->n_clusters];
    b->n_paths = cl->path_offset[cl->n_clusters];
    b->nnz = cl->row_ptr[b->n_reads];

    b->h_read_offset.assign(cl->read_offset, cl->read_offset + cl->n_clusters + 1);
    b->h_path_offset.assign(cl->path_offset, cl->path_offset + cl->n_clusters + 1);
    b->h_cluster_nnz.resize(cl->n_clusters);

    for (uint32_t c = 0; c < cl->n_clusters; ++c) {
        const uint64_t lo = cl->row_ptr[cl->read_offset[c]];
        const uint64_t hi = cl->row_ptr[cl->read_offset[c + 1]];
        if (hi < lo) {
            delete b;
            return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "row_ptr must be non-decreasing");
        }
        b->h_cluster_nnz[c] = hi - lo;
    }

    // Largest first, like the reference's schedule (rpvg:src/main.cpp:827): the hardware block scheduler then hands the
    // long-running clusters out first.
    std::vector<uint32_t> order(cl->n_clusters);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) {
        const uint64_t cx = b->h_cluster_nnz[x] + (b->h_read_offset[x + 1] - b->h_read_offset[x]);
        const uint64_t cy = b->h_cluster_nnz[y] + (b->h_read_offset[y + 1] - b->h_read_offset[y]);
        return cx > cy;
    });

    const uint64_t C = b->n_clusters, R = b->n_reads, P = b->n_paths, Z = b->nnz;

    b->sort_bits = 1;
    while ((uint64_t(1) << b->sort_bits) < P + 1 && b->sort_bits < 32) { b->sort_bits++; }

    cub::DeviceRadixSort::SortPairs(nullptr, b->sort_temp_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                    static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<int>(Z), 0, b->sort_bits, ctx->stream);

    size_t bytes = 0;
    bytes += 2 * padded<uint64_t>(C + 1) + padded<uint64_t>(R + 1) + padded<uint32_t>(Z) + padded<double>(Z) + padded<double>(R) + padded<uint32_t>(R) + padded<uint32_t>(C);
    bytes += padded<double>(Z) + padded<double>(R) + 5 * padded<uint32_t>(Z) + padded<uint64_t>(P + 1) + padded<uint32_t>(Z) + padded<double>(Z);
    bytes += ((b->sort_temp_bytes + 255) & ~size_t(255));
    bytes += padded<double>(P + C) + padded<double>(R) + padded<double>(P) + 2 * padded<double>(C) + padded<uint32_t>(C) + padded<uint32_t>(1);
    bytes += 4096;

    cudaError_t err = cudaMallocAsync(reinterpret_cast<void **>(&b->pool.base), bytes, ctx->stream);
    if (err != cudaSuccess) {
        delete b;
        return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("cudaMallocAsync(") + std::to_string(bytes) + "): " + cudaGetErrorString(err));
    }
    b->pool.size = bytes;

    auto & pool = b->pool;
    b->d_read_offset = pool.take<uint64_t>(C + 1);
    b->d_path_offset = pool.take<uint64_t>(C + 1);
    b->d_row_ptr = pool.take<uint64_t>(R + 1);
    b->d_col = pool.take<uint32_t>(Z);
    b->d_prob = pool.take<double>(Z);
    b->d_noise = pool.take<double>(R);
    b->d_count = pool.take<uint32_t>(R);
    b->d_order = pool.take<uint32_t>(C);
    b->d_val = pool.take<double>(Z);
    b->d_cnt = pool.take<double>(R);
    b->d_entry_row = pool.take<uint32_t>(Z);
    b->d_key_in = pool.take<uint32_t>(Z);
    b->d_key_out = pool.take<uint32_t>(Z);
    b->d_idx_in = pool.take<uint32_t>(Z);
    b->d_idx_out = pool.take<uint32_t>(Z);
    b->d_col_ptr = pool.take<uint64_t>(P + 1);
    b->d_csc_row = pool.take<uint32_t>(Z);
    b->d_csc_val = pool.take<double>(Z);
    b->d_sort_temp = pool.take<char>(b->sort_temp_bytes);
    b->d_abund = pool.take<double>(P + C);
    b->d_w = pool.take<double>(R);
    b->d_abundances = pool.take<double>(P);
    b->d_noise_count = pool.take<double>(C);
    b->d_total_count = pool.take<double>(C);
    b->d_iterations = pool.take<uint32_t>(C);
    b->d_error = pool.take<uint32_t>(1);

    if (pool.used > pool.size) {
        cudaFreeAsync(pool.base, ctx->stream);
        delete b;
        return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "internal: device pool overflow");
    }

    cudaStream_t s = ctx->stream;

    #define RPVG_UPLOAD(dst, src, n, T)                                                                             \
        if ((n) > 0) {                                                                                              \
            err = cudaMemcpyAsync((dst), (src), (n) * sizeof(T), cudaMemcpyHostToDevice, s);                        \
            if (err != cudaSuccess) {                                                                               \
                cudaFreeAsync(pool.base, s);                                                                        \
                delete b;                                                                                           \
                return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("H2D " #src ": ") + cudaGetErrorString(err));  \
            }                                                                                                       \
        }

    RPVG_UPLOAD(b->d_read_offset, cl->read_offset, C + 1, uint64_t)
    RPVG_UPLOAD(b->d_path_offset, cl->path_offset, C + 1, uint64_t)
    RPVG_UPLOAD(b->d_row_ptr, cl->row_ptr, R + 1, uint64_t)
    RPVG_UPLOAD(b->d_col, cl->col, Z, uint32_t)
    RPVG_UPLOAD(b->d_prob, cl->prob, Z, double)
    RPVG_UPLOAD(b->d_noise, cl->noise, R, double)
    RPVG_UPLOAD(b->d_count, cl->count, R, uint32_t)
    // order is a small temporary host vector: the copy must finish before it goes out of scope
    RPVG_UPLOAD(b->d_order, order.data(), C, uint32_t)
    #undef RPVG_UPLOAD

    err = cudaStreamSynchronize(s);
    if (err != cudaSuccess) {
        cudaFreeAsync(pool.base, s);
        delete b;
        return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("upload synchronize: ") + cudaGetErrorString(err));
    }

    *out_batch = b;
    return RPVG_B200_OK;
}

int rpvg_b200_batch_em(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, const rpvg_b200_em_params * params) {

    if (!ctx || !b || !params) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (params->min_em_conv_its == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "min_em_conv_its must be > 0"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    if (!b->ev_start) {
        RPVG_CUDA(ctx, cudaEventCreate(&b->ev_start));
        RPVG_CUDA(ctx, cudaEventCreate(&b->ev_built));
        RPVG_CUDA(ctx, cudaEventCreate(&b->ev_done));
    }

    RPVG_CUDA(ctx, cudaEventRecord(b->ev_start, s));
    RPVG_CUDA(ctx, cudaMemsetAsync(b->d_error, 0, sizeof(uint32_t), s));

    if (b->n_reads > 0) {

        const unsigned grid = static_cast<unsigned>((b->n_reads + 255) / 256);
        buildRowsKernel<<<grid, 256, 0, s>>>(b->n_clusters, b->n_reads, b->d_read_offset, b->d_path_offset, b->d_row_ptr, b->d_col, b->d_prob,
                                             b->d_noise, b->d_count, b->d_val, b->d_cnt, b->d_entry_row, b->d_key_in, b->d_idx_in, b->d_error);
        RPVG_LAUNCHED(ctx);
    }

    if (b->nnz > 0) {

        size_t temp_bytes = b->sort_temp_bytes;
        RPVG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(b->d_sort_temp, temp_bytes, b->d_key_in, b->d_key_out, b->d_idx_in, b->d_idx_out,
                                                       static_cast<int>(b->nnz), 0, b->sort_bits, s));
        ctx->launches += static_cast<uint64_t>((b->sort_bits + 7) / 8) * 3;   // library launches, counted approximately
    }

    {
        const unsigned grid = static_cast<unsigned>((b->n_paths + 1 + 255) / 256);
        columnPointerKernel<<<grid, 256, 0, s>>>(b->n_paths, b->nnz, b->d_key_out, b->d_col_ptr);
        RPVG_LAUNCHED(ctx);
    }

    if (b->nnz > 0) {

        const unsigned grid = static_cast<unsigned>((b->nnz + 255) / 256);
        gatherColumnsKernel<<<grid, 256, 0, s>>>(b->nnz, b->d_idx_out, b->d_entry_row, b->d_val, b->d_csc_row, b->d_csc_val);
        RPVG_LAUNCHED(ctx);
    }

    EmArgs args;
    args.order = b->d_order;
    args.read_offset = b->d_read_offset;
    args.path_offset = b->d_path_offset;
    args.row_ptr = b->d_row_ptr;
    args.col = b->d_col;
    args.val = b->d_val;
    args.noise = b->d_noise;
    args.cnt = b->d_cnt;
    args.col_ptr = b->d_col_ptr;
    args.csc_row = b->d_csc_row;
    args.csc_val = b->d_csc_val;
    args.abund = b->d_abund;
    args.w = b->d_w;
    args.abundances = b->d_abundances;
    args.noise_count = b->d_noise_count;
    args.total_count = b->d_total_count;
    args.iterations = b->d_iterations;
    args.params = *params;

    RPVG_CUDA(ctx, cudaEventRecord(b->ev_built, s));

    emClusterKernel<<<b->n_clusters, kEmThreads, 0, s>>>(args);
    RPVG_LAUNCHED(ctx);

    RPVG_CUDA(ctx, cudaEventRecord(b->ev_done, s));

    b->have_results = true;
    return RPVG_B200_OK;
}

int rpvg_b200_batch_download(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, double * abundances, double * noise_count, double * total_count, uint32_t * iterations) {

    if (!ctx || !b) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (!b->have_results) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_batch_em has not run on this batch"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    b->h_iterations.resize(b->n_clusters);
    uint32_t dev_error = 0;

    if (abundances) { RPVG_CUDA(ctx, cudaMemcpyAsync(abundances, b->d_abundances, b->n_paths * sizeof(double), cudaMemcpyDeviceToHost, s)); }
    if (noise_count) { RPVG_CUDA(ctx, cudaMemcpyAsync(noise_count, b->d_noise_count, b->n_clusters * sizeof(double), cudaMemcpyDeviceToHost, s)); }
    if (total_count) { RPVG_CUDA(ctx, cudaMemcpyAsync(total_count, b->d_total_count, b->n_clusters * sizeof(double), cudaMemcpyDeviceToHost, s)); }
    RPVG_CUDA(ctx, cudaMemcpyAsync(b->h_iterations.data(), b->d_iterations, b->n_clusters * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    RPVG_CUDA(ctx, cudaMemcpyAsync(&dev_error, b->d_error, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    RPVG_CUDA(ctx, cudaStreamSynchronize(s));

    if (iterations) { std::memcpy(iterations, b->h_iterations.data(), b->n_clusters * sizeof(uint32_t)); }

    if (dev_error) {

        std::string what = "malformed cluster arena:";
        if (dev_error & RPVG_DEV_ERR_COL_RANGE) { what += " path index out of range;"; }
        if (dev_error & RPVG_DEV_ERR_COL_ORDER) { what += " entries of a read class not strictly ascending by path index;"; }
        if (dev_error & RPVG_DEV_ERR_VALUE) { what += " negative or non-finite probability;"; }
        return rpvgFail(ctx, RPVG_B200_ERR_INVALID, what);
    }

    return RPVG_B200_OK;
}

int rpvg_b200_batch_em_work(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, double * nnz_iterations, double * algorithmic_bytes) {

    if (!ctx || !b) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (b->h_iterations.size() != b->n_clusters) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "results have not been downloaded"); }

    double work = 0, bytes = 0;

    for (uint32_t c = 0; c < b->n_clusters; ++c) {

        const double its = b->h_iterations[c];
        const double M = static_cast<double>(b->h_read_offset[c + 1] - b->h_read_offset[c]);
        const double N = static_cast<double>(b->h_path_offset[c + 1] - b->h_path_offset[c]);
        const double nnz_path = static_cast<double>(b->h_cluster_nnz[c]);

        work += its * (nnz_path + M);
        bytes += its * (12 * nnz_path + 20 * M + 16 * (N + 1));
    }

    if (nnz_iterations) { *nnz_iterations = work; }
    if (algorithmic_bytes) { *algorithmic_bytes = bytes; }
    return RPVG_B200_OK;
}

int rpvg_b200_batch_em_timings(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, float * build_ms, float * em_ms) {

    if (!ctx || !b) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (!b->ev_start) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_batch_em has not run on this batch"); }

    RPVG_CUDA(ctx, cudaEventSynchronize(b->ev_done));
    if (build_ms) { RPVG_CUDA(ctx, cudaEventElapsedTime(build_ms, b->ev_start, b->ev_built)); }
    if (em_ms) { RPVG_CUDA(ctx, cudaEventElapsedTime(em_ms, b->ev_built, b->ev_done)); }
    return RPVG_B200_OK;
}

void rpvg_b200_batch_free(rpvg_b200_ctx * ctx, rpvg_b200_batch * b) {

    if (!b) { return; }
    if (b->ev_start) {
        cudaEventDestroy(b->ev_start);
        cudaEventDestroy(b->ev_built);
        cudaEventDestroy(b->ev_done);
    }
    if (ctx && b->pool.base) {
        cudaSetDevice(ctx->device);
        cudaFreeAsync(b->pool.base, ctx->stream);
    }
    delete b;
}

int rpvg_b200_em_abundances(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_em_params * params,
                            double * abundances, double * noise_count, double * total_count, uint32_t * iterations) {

    rpvg_b200_batch * batch = nullptr;

    int status = rpvg_b200_batch_upload(ctx, clusters, &batch);
    if (status != RPVG_B200_OK) { return status; }

    status = rpvg_b200_batch_em(ctx, batch, params);
    if (status == RPVG_B200_OK) { status = rpvg_b200_batch_download(ctx, batch, abundances, noise_count, total_count, iterations); }

    rpvg_b200_batch_free(ctx, batch);
    return status;
}

}
