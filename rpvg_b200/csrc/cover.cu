// rpvg_b200_path_cover — the greedy weighted minimum path cover of the strains model, for a batch of clusters.
//
// Reference semantics restated (never its code): MinimumPathAbundanceEstimator::estimate builds
//   cover[i][j]  = read class i has path j,
//   count_i      = read count, 0 when noise_i == 1 under doubleCompare,
//   weight_j     = - sum_i log(p_ij) * count_i                      (rpvg:src/path_abundance_estimator.cpp:229-251)
// and weightedMinimumPathCover (:297-340) picks, while any uncovered count is > 0, the FIRST path with the strictly
// largest (sum of uncovered counts it covers) / weight_j, then zeroes the read classes it covers; one path -> {0}.
//
// Device: entries sorted by (global path, read class) with a stable radix sort, so that a column's weight is summed in
// read-class order like the reference does; one CTA per cluster runs the pick loop on uncovered counts that are kept up
// to date per column (see coverKernel): a block arg-max that prefers the lower index among equal scores, then the rows of
// the picked column leave the counts of every column they are in. Sums of integers: exact, hence deterministic.

#include <cfloat>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int kCoverThreads = 256;

__global__ void coverRowsKernel(uint32_t n_clusters, uint64_t n_reads, const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ path_offset,
                                const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col, const double * __restrict__ noise,
                                const uint32_t * __restrict__ count, uint32_t * __restrict__ key, uint32_t * __restrict__ idx, uint32_t * __restrict__ entry_row,
                                double * __restrict__ cnt, uint32_t * __restrict__ error) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_reads) { return; }

    const uint32_t c = clusterOf(read_offset, n_clusters, r);
    const uint64_t base = path_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - base;

    for (uint64_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) {

        uint32_t j = col[e];
        if (j >= num_paths) {
            atomicOr(error, RPVG_DEV_ERR_COL_RANGE);
            j = 0;
        }
        key[e] = static_cast<uint32_t>(base + j);
        idx[e] = static_cast<uint32_t>(e);
        entry_row[e] = static_cast<uint32_t>(r);
    }

    // doubleCompare(noise, 1): rpvg:src/utils.hpp:87-93
    const double nz = noise[r];
    const bool is_noise = (nz == 1.0) || (fabs(nz - 1.0) < fabs(fmin(nz, 1.0)) * (DBL_EPSILON * 100));
    cnt[r] = is_noise ? 0.0 : static_cast<double>(count[r]);
}

__global__ void coverColumnPointerKernel(uint64_t n_paths, uint64_t nnz, const uint32_t * __restrict__ sorted_key, uint64_t * __restrict__ col_ptr) {

    const uint64_t j = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (j > n_paths) { return; }

    uint64_t lo = 0, hi = nnz;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (__ldg(sorted_key + mid) >= j) { hi = mid; } else { lo = mid + 1; }
    }
    col_ptr[j] = lo;
}

// thread per path: weight in read-class order; rows of the column
__global__ void coverWeightsKernel(uint64_t n_paths, const uint64_t * __restrict__ col_ptr, const uint32_t * __restrict__ sorted_idx, const uint32_t * __restrict__ entry_row,
                                   const double * __restrict__ prob, const double * __restrict__ cnt, uint32_t * __restrict__ csc_row, double * __restrict__ weight) {

    const uint64_t j = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (j >= n_paths) { return; }

    double w = 0;
    for (uint64_t k = col_ptr[j]; k < col_ptr[j + 1]; ++k) {
        const uint32_t e = sorted_idx[k];
        const uint32_t r = entry_row[e];
        csc_row[k] = r;
        w += log(prob[e]) * cnt[r];
    }
    weight[j] = w * -1;
}

// One CTA per cluster. The reference recomputes every column's uncovered count for every pick (O(picks x entries)); here
// the counts are kept up to date instead: a read class is zeroed exactly once, and when it is, its count is taken off
// every column that holds it (the entries of its row). All sums are sums of integers below 2^53 — exact in fp64 whatever
// the order, atomics included — so the scores, and with them the picks, are the reference's. Work: O(entries + picks x N).
// The picks of a cluster are sequential and each is a scan over its paths: the clusters with many paths (`list`, or all
// when it is NULL) get a CTA of 1024 threads, the others one of 256.
template <int THREADS>
__global__ void __launch_bounds__(THREADS) coverKernel(const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ path_offset,
                                                       const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col,
                                                       const uint64_t * __restrict__ col_ptr, const uint32_t * __restrict__ csc_row,
                                                       const double * __restrict__ weight, double * uncovered, double * covered, uint8_t * __restrict__ in_cover,
                                                       const uint32_t * __restrict__ error, const uint32_t * __restrict__ list) {

    constexpr int kCoverThreads = THREADS;   // (shadows the small kernels' constant)

    __shared__ double best_score[kCoverThreads / 32];
    __shared__ uint32_t best_index[kCoverThreads / 32];
    __shared__ double red[32];
    __shared__ int pick_shared;

    if (*error) { return; }

    const uint32_t c = list ? list[blockIdx.x] : blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int num_warps = kCoverThreads / 32;

    const uint64_t r0 = read_offset[c], r1 = read_offset[c + 1];
    const uint64_t p0 = path_offset[c];
    const uint32_t N = static_cast<uint32_t>(path_offset[c + 1] - p0);

    for (uint32_t j = tid; j < N; j += kCoverThreads) { in_cover[p0 + j] = 0; }

    if (r1 == r0) { return; }   // the reference does not get here without read classes

    if (N == 1) {
        if (tid == 0) { in_cover[p0] = 1; }
        return;
    }

    // uncovered count of every column, and the total still uncovered
    for (uint32_t j = warp; j < N; j += num_warps) {
        double sum = 0;
        for (uint64_t k = col_ptr[p0 + j] + lane; k < col_ptr[p0 + j + 1]; k += 32) { sum += uncovered[csc_row[k]]; }
        sum = warpSum(sum);
        if (lane == 0) { covered[p0 + j] = sum; }
    }

    double mine = 0;
    for (uint64_t r = r0 + tid; r < r1; r += kCoverThreads) { mine += uncovered[r]; }
    double remaining = blockSum(mine, red);

    __syncthreads();

    for (uint32_t pick_round = 0; pick_round < N; ++pick_round) {

        // while (uncovered_read_counts.maxCoeff() > 0): counts are non-negative integers
        if (!(remaining > 0)) { break; }

        double score = 0;            // strictly-greater-than-0 is what the reference's scan starts from
        uint32_t index = 0xffffffffu;

        for (uint32_t j = tid; j < N; j += kCoverThreads) {   // ascending j per thread: its first maximum stays
            const double count = covered[p0 + j];
            if (!(count > 0)) { continue; }   // (0 / weight is never above the score the scan starts from: most paths, soon)
            const double s = count / weight[p0 + j];
            if (s > score) {
                score = s;
                index = j;
            }
        }

        // first strict maximum of the block: larger score, or the same score at a lower index
        #pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double other_score = __shfl_xor_sync(0xffffffffu, score, off);
            const uint32_t other_index = __shfl_xor_sync(0xffffffffu, index, off);
            if (other_index != 0xffffffffu && (index == 0xffffffffu || other_score > score || (other_score == score && other_index < index))) {
                score = other_score;
                index = other_index;
            }
        }

        if (lane == 0) {
            best_score[warp] = score;
            best_index[warp] = index;
        }
        __syncthreads();

        if (warp == 0) {

            double s = lane < num_warps ? best_score[lane] : 0.0;
            uint32_t best = lane < num_warps ? best_index[lane] : 0xffffffffu;

            #pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const double other_score = __shfl_xor_sync(0xffffffffu, s, off);
                const uint32_t other_index = __shfl_xor_sync(0xffffffffu, best, off);
                if (other_index != 0xffffffffu && (best == 0xffffffffu || other_score > s || (other_score == s && other_index < best))) {
                    s = other_score;
                    best = other_index;
                }
            }

            if (lane == 0) { pick_shared = static_cast<int>(best); }
        }
        __syncthreads();

        const uint32_t pick = static_cast<uint32_t>(pick_shared);
        if (pick == 0xffffffffu) { break; }   // the reference asserts here (no positive score)

        if (tid == 0) { in_cover[p0 + pick] = 1; }

        // the read classes the pick covers leave every column that holds them
        double taken = 0;

        for (uint64_t k = col_ptr[p0 + pick] + tid; k < col_ptr[p0 + pick + 1]; k += kCoverThreads) {

            const uint64_t r = csc_row[k];   // global read class (rows of a column are distinct: no two threads meet on one)
            const double cnt = uncovered[r];

            if (cnt > 0) {
                uncovered[r] = 0;
                taken += cnt;
                for (uint64_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) { atomicAdd(&covered[p0 + col[e]], -cnt); }
            }
        }

        remaining -= blockSum(taken, red);   // (barriers inside: the updates are done before the next scan)
        __syncthreads();
    }
}

}

extern "C" int rpvg_b200_path_cover(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, uint8_t * in_cover) {

    if (!ctx || !cl || !in_cover) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (cl->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no clusters"); }
    { const int shape = rpvgCheckClusters(ctx, cl); if (shape != RPVG_B200_OK) { return shape; } }

    const uint32_t C = cl->n_clusters;
    const uint64_t R = cl->read_offset[C];
    const uint64_t P = cl->path_offset[C];
    const uint64_t Z = R > 0 ? cl->row_ptr[R] : 0;

    if (P >= 0xffffffffull || Z >= 0xffffffffull || R >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "batch too large (paths, read classes, entries below 2^32)"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    int sort_bits = 1;
    while ((uint64_t(1) << sort_bits) < P + 1 && sort_bits < 32) { sort_bits++; }

    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                    static_cast<uint32_t *>(nullptr), static_cast<int>(Z), 0, sort_bits, s);

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 2 * pad(8 * (C + 1)) + pad(8 * (R + 1)) + pad(4 * Z) + pad(8 * Z) + pad(8 * R) + pad(4 * R) + 6 * pad(4 * Z) + pad(8 * R) + pad(8 * (P + 1)) +
                         2 * pad(8 * P) + pad(P) + pad(sort_bytes) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_read_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_row_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_col = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_prob = reinterpret_cast<double *>(take(8 * Z));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_count = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_key_in = reinterpret_cast<uint32_t *>(take(4 * Z));
    uint32_t * d_key_out = reinterpret_cast<uint32_t *>(take(4 * Z));
    uint32_t * d_idx_in = reinterpret_cast<uint32_t *>(take(4 * Z));
    uint32_t * d_idx_out = reinterpret_cast<uint32_t *>(take(4 * Z));
    uint32_t * d_entry_row = reinterpret_cast<uint32_t *>(take(4 * Z));
    uint32_t * d_csc_row = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_cnt = reinterpret_cast<double *>(take(8 * R));
    uint64_t * d_col_ptr = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    double * d_weight = reinterpret_cast<double *>(take(8 * P));
    double * d_covered = reinterpret_cast<double *>(take(8 * P));
    uint8_t * d_in_cover = reinterpret_cast<uint8_t *>(take(P));
    void * d_sort_temp = take(sort_bytes);
    uint32_t * d_error = reinterpret_cast<uint32_t *>(take(4));

    cudaError_t err = cudaSuccess;
    int status = RPVG_B200_OK;

    #define RPVG_STEP(call) if (err == cudaSuccess) { err = (call); }
    #define RPVG_H2D(dst, src, n) if (err == cudaSuccess && (n) > 0) { err = cudaMemcpyAsync((dst), (src), (n), cudaMemcpyHostToDevice, s); }

    RPVG_H2D(d_read_offset, cl->read_offset, 8 * (C + 1))
    RPVG_H2D(d_path_offset, cl->path_offset, 8 * (C + 1))
    RPVG_H2D(d_row_ptr, cl->row_ptr, 8 * (R + 1))
    RPVG_H2D(d_col, cl->col, 4 * Z)
    RPVG_H2D(d_prob, cl->prob, 8 * Z)
    RPVG_H2D(d_noise, cl->noise, 8 * R)
    RPVG_H2D(d_count, cl->count, 4 * R)
    RPVG_STEP(cudaMemsetAsync(d_error, 0, 4, s))

    if (err == cudaSuccess && R > 0) {
        coverRowsKernel<<<static_cast<unsigned>((R + 255) / 256), 256, 0, s>>>(C, R, d_read_offset, d_path_offset, d_row_ptr, d_col, d_noise, d_count, d_key_in, d_idx_in,
                                                                               d_entry_row, d_cnt, d_error);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess && Z > 0) {
        size_t temp = sort_bytes;
        err = cub::DeviceRadixSort::SortPairs(d_sort_temp, temp, d_key_in, d_key_out, d_idx_in, d_idx_out, static_cast<int>(Z), 0, sort_bits, s);
        ctx->launches += 2 + (sort_bits + 7) / 8;
    }

    if (err == cudaSuccess) {
        coverColumnPointerKernel<<<static_cast<unsigned>((P + 1 + 255) / 256), 256, 0, s>>>(P, Z, d_key_out, d_col_ptr);
        coverWeightsKernel<<<static_cast<unsigned>((P + 255) / 256), 256, 0, s>>>(P, d_col_ptr, d_idx_out, d_entry_row, d_prob, d_cnt, d_csc_row, d_weight);
        // clusters of many paths on CTAs of 1024 threads (first: they take longest), the others on CTAs of 256
        std::vector<uint32_t> wide, narrow;
        for (uint32_t c = 0; c < C; ++c) { (cl->path_offset[c + 1] - cl->path_offset[c] >= 512 ? wide : narrow).push_back(c); }

        uint32_t * d_list = nullptr;
        err = cudaMallocFromPoolAsync(reinterpret_cast<void **>(&d_list), 4 * static_cast<size_t>(C), ctx->pool, s);
        if (err == cudaSuccess && !wide.empty()) { err = cudaMemcpyAsync(d_list, wide.data(), 4 * wide.size(), cudaMemcpyHostToDevice, s); }
        if (err == cudaSuccess && !narrow.empty()) { err = cudaMemcpyAsync(d_list + wide.size(), narrow.data(), 4 * narrow.size(), cudaMemcpyHostToDevice, s); }

        // (the narrow ones on a side stream beside the wide ones: a few CTAs of 1024 threads do not fill the GPU)
        cudaStream_t side = (!wide.empty() && !narrow.empty() && !ctx->side_streams.empty()) ? ctx->side_streams[0] : s;

        if (err == cudaSuccess && side != s) {
            cudaEventRecord(ctx->fork_event, s);
            cudaStreamWaitEvent(side, ctx->fork_event, 0);
        }
        if (err == cudaSuccess && !wide.empty()) {
            coverKernel<1024><<<static_cast<unsigned>(wide.size()), 1024, 0, s>>>(d_read_offset, d_path_offset, d_row_ptr, d_col, d_col_ptr, d_csc_row, d_weight, d_cnt, d_covered, d_in_cover,
                                                                                  d_error, d_list);
            ctx->launches++;
        }
        if (err == cudaSuccess && !narrow.empty()) {
            coverKernel<kCoverThreads><<<static_cast<unsigned>(narrow.size()), kCoverThreads, 0, side>>>(d_read_offset, d_path_offset, d_row_ptr, d_col, d_col_ptr, d_csc_row, d_weight, d_cnt,
                                                                                                         d_covered, d_in_cover, d_error, d_list + wide.size());
        }
        if (err == cudaSuccess && side != s) {
            cudaEventRecord(ctx->join_events[0], side);
            cudaStreamWaitEvent(s, ctx->join_events[0], 0);
        }
        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }   // (the lists are host vectors)
        if (d_list) { cudaFreeAsync(d_list, s); }
        ctx->launches += 3;
        err = cudaGetLastError();
    }

    uint32_t dev_error = 0;
    RPVG_STEP(cudaMemcpyAsync(in_cover, d_in_cover, P, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaMemcpyAsync(&dev_error, d_error, 4, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaStreamSynchronize(s))

    #undef RPVG_STEP
    #undef RPVG_H2D

    if (err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_path_cover: ") + cudaGetErrorString(err)); }
    else if (dev_error) { status = rpvgFail(ctx, RPVG_B200_ERR_INVALID, "malformed cluster arena: path index out of range"); }

    cudaFreeAsync(base, s);
    return status;
}


// ---- weightedMinimumPathCover itself (rpvg:src/path_abundance_estimator.cpp:297-340) on a caller-made cover ---------------
namespace {

// one thread per read class: keys of its cover entries (the read classes come with their counts as they are)
__global__ void coverGivenRowsKernel(uint64_t n_reads, const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col, const uint32_t n_paths,
                                     uint32_t * __restrict__ key, uint32_t * __restrict__ idx, uint32_t * __restrict__ entry_row, uint32_t * __restrict__ error) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_reads) { return; }

    for (uint64_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) {
        uint32_t j = col[e];
        if (j >= n_paths) {
            atomicOr(error, RPVG_DEV_ERR_COL_RANGE);
            j = 0;
        }
        key[e] = j;
        idx[e] = static_cast<uint32_t>(e);
        entry_row[e] = static_cast<uint32_t>(r);
    }
}

__global__ void coverGivenColumnsKernel(uint64_t nnz, const uint32_t * __restrict__ sorted_idx, const uint32_t * __restrict__ entry_row, uint32_t * __restrict__ csc_row) {

    const uint64_t k = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (k < nnz) { csc_row[k] = entry_row[sorted_idx[k]]; }
}

}

extern "C" int rpvg_b200_weighted_path_cover(rpvg_b200_ctx * ctx, uint32_t n_reads, uint32_t n_paths, const uint64_t * row_ptr, const uint32_t * col,
                                             const double * read_counts, const double * path_weights, uint8_t * in_cover) {

    if (!ctx || !row_ptr || !read_counts || !path_weights || !in_cover) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_weighted_path_cover: NULL argument"); }
    if (n_paths == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_weighted_path_cover: no paths"); }

    if (n_reads == 0) {   // nothing uncovered: the reference's loop does not run (one path -> {0} all the same)
        for (uint32_t j = 0; j < n_paths; ++j) { in_cover[j] = (n_paths == 1) ? 1 : 0; }
        return RPVG_B200_OK;
    }

    const uint64_t R = n_reads, P = n_paths, Z = row_ptr[R];
    if (row_ptr[0] != 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_weighted_path_cover: row_ptr must start at 0"); }
    if (Z >= 0x7fffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_weighted_path_cover: too many entries"); }
    if (Z > 0 && !col) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_weighted_path_cover: col is NULL"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    int sort_bits = 1;
    while ((uint64_t(1) << sort_bits) < P + 1 && sort_bits < 32) { sort_bits++; }

    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                    static_cast<uint32_t *>(nullptr), static_cast<int>(Z), 0, sort_bits, s);

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 2 * pad(16) + pad(8 * (R + 1)) + 7 * pad(4 * (Z + 1)) + pad(8 * R) + pad(8 * (P + 1)) + 2 * pad(8 * P) + pad(P) + pad(sort_bytes) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_read_offset = reinterpret_cast<uint64_t *>(take(16));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(16));
    uint64_t * d_row_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_col = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    uint32_t * d_key_in = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    uint32_t * d_key_out = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    uint32_t * d_idx_in = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    uint32_t * d_idx_out = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    uint32_t * d_entry_row = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    uint32_t * d_csc_row = reinterpret_cast<uint32_t *>(take(4 * (Z + 1)));
    double * d_cnt = reinterpret_cast<double *>(take(8 * R));
    uint64_t * d_col_ptr = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    double * d_weight = reinterpret_cast<double *>(take(8 * P));
    double * d_covered = reinterpret_cast<double *>(take(8 * P));
    uint8_t * d_in_cover = reinterpret_cast<uint8_t *>(take(P));
    void * d_sort_temp = take(sort_bytes);
    uint32_t * d_error = reinterpret_cast<uint32_t *>(take(4));

    const uint64_t read_offset[2] = {0, R}, path_offset[2] = {0, P};

    cudaError_t err = cudaMemcpyAsync(d_read_offset, read_offset, 16, cudaMemcpyHostToDevice, s);
    #define RPVG_STEP(call) if (err == cudaSuccess) { err = (call); }
    RPVG_STEP(cudaMemcpyAsync(d_path_offset, path_offset, 16, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_row_ptr, row_ptr, 8 * (R + 1), cudaMemcpyHostToDevice, s))
    if (Z > 0) { RPVG_STEP(cudaMemcpyAsync(d_col, col, 4 * Z, cudaMemcpyHostToDevice, s)) }
    RPVG_STEP(cudaMemcpyAsync(d_cnt, read_counts, 8 * R, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_weight, path_weights, 8 * P, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemsetAsync(d_error, 0, 4, s))
    RPVG_STEP(cudaStreamSynchronize(s))   // (the two offset pairs are on this frame)

    if (err == cudaSuccess) {
        coverGivenRowsKernel<<<static_cast<unsigned>((R + 255) / 256), 256, 0, s>>>(R, d_row_ptr, d_col, n_paths, d_key_in, d_idx_in, d_entry_row, d_error);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess && Z > 0) {
        size_t temp = sort_bytes;
        err = cub::DeviceRadixSort::SortPairs(d_sort_temp, temp, d_key_in, d_key_out, d_idx_in, d_idx_out, static_cast<int>(Z), 0, sort_bits, s);
        ctx->launches += 2 + (sort_bits + 7) / 8;
    }

    if (err == cudaSuccess) {
        coverColumnPointerKernel<<<static_cast<unsigned>((P + 1 + 255) / 256), 256, 0, s>>>(P, Z, d_key_out, d_col_ptr);
        if (Z > 0) { coverGivenColumnsKernel<<<static_cast<unsigned>((Z + 255) / 256), 256, 0, s>>>(Z, d_idx_out, d_entry_row, d_csc_row); }
        if (n_paths >= 512) {
            coverKernel<1024><<<1, 1024, 0, s>>>(d_read_offset, d_path_offset, d_row_ptr, d_col, d_col_ptr, d_csc_row, d_weight, d_cnt, d_covered, d_in_cover, d_error, nullptr);
        } else {
            coverKernel<kCoverThreads><<<1, kCoverThreads, 0, s>>>(d_read_offset, d_path_offset, d_row_ptr, d_col, d_col_ptr, d_csc_row, d_weight, d_cnt, d_covered, d_in_cover, d_error, nullptr);
        }
        ctx->launches += 3;
        err = cudaGetLastError();
    }

    uint32_t dev_error = 0;
    RPVG_STEP(cudaMemcpyAsync(in_cover, d_in_cover, P, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaMemcpyAsync(&dev_error, d_error, 4, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaStreamSynchronize(s))
    #undef RPVG_STEP

    cudaFreeAsync(base, s);

    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_weighted_path_cover: ") + cudaGetErrorString(err)); }
    if (dev_error) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_weighted_path_cover: path index out of range"); }
    return RPVG_B200_OK;
}
