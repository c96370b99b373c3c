// Bounded haplotype-pair posteriors (rpvg:src/path_estimator.cpp:379-473), the part that is O(#pairs): for every
// cluster of a resident batch (grouped.cuh) all pairs (i, j >= i) of its groups, taken in the order the caller gives
// (the reference orders the groups by marginal posterior), get
//   ll = sum_r count_r log(noise_r + (col_i + col_j) / 2) + log_freq_i + log_freq_j + log(numPermutations)
// and the reference then walks them once, skipping a pair when ll - (largest ll so far) < log(min_rel_likelihood).
// A skipped pair is smaller than the running maximum, so it never changes it: the running maximum is the prefix maximum
// over all pairs, and "kept" is a function of the pair's own value and the prefix maximum before it. The kernels here
// compute the values, the exclusive prefix maximum per cluster and hand back only the pairs that pass the test with a
// safety margin; the caller replays the reference's exact walk on those few (any pair dropped here fails the exact test
// too and cannot move the maximum), which gives the same list and the same numbers as walking all pairs.
//
// One CTA per cluster: pairs of a cluster are few thousand; a warp per pair for the sum over read classes (as in
// posterior.cu), then a block-wide scan. fp64 log bound.

#include <cfloat>
#include <string>
#include <vector>

#include "common.cuh"
#include "grouped.cuh"

namespace {

constexpr int kBoundedThreads = 256;

struct BoundedArgs {

    const double * matrix;             // the resident batch (grouped.cuh)
    const double * noise;
    const double * counts;
    const uint64_t * row_offset;
    const uint64_t * group_offset;
    const uint64_t * matrix_offset;
    const uint32_t * n_classes;
    const uint64_t * pair_offset;      // [n_clusters + 1] pairs of each cluster = G (G + 1) / 2
    const uint32_t * order;            // per cluster (at group_offset): group ids in the caller's order
    const double * log_freq;           // per group (at group_offset), indexed by group id
    double * pair_ll;                  // [n_pairs] scratch
    double min_log_diff;               // log(min_rel_likelihood) - margin
    uint32_t * n_candidates;           // [n_clusters]
    uint32_t * cand_pair;              // [n_pairs] candidates of a cluster, compacted at pair_offset[c]
    double * cand_loglik;              // the raw sum over read classes of each candidate
};

// pair index -> (i, j), pairs ordered (0,0) (0,1) .. (0,G-1) (1,1) ..
__device__ __forceinline__ void pairOfIndex(const uint32_t G, const uint32_t index, uint32_t * i_out, uint32_t * j_out) {

    // row i starts at i * G - i (i - 1) / 2
    uint32_t i = static_cast<uint32_t>(floor((2.0 * G + 1.0 - sqrt((2.0 * G + 1.0) * (2.0 * G + 1.0) - 8.0 * index)) / 2.0));
    auto row_start = [&](uint32_t r) { return r * G - (r * (r - 1)) / 2; };
    while (i > 0 && row_start(i) > index) { --i; }
    while (i + 1 < G && row_start(i + 1) <= index) { ++i; }
    *i_out = i;
    *j_out = i + (index - row_start(i));
}

__global__ void __launch_bounds__(kBoundedThreads) boundedPairsKernel(const BoundedArgs args) {

    __shared__ double warp_max[kBoundedThreads / 32];
    __shared__ double carry;
    __shared__ uint32_t warp_count[kBoundedThreads / 32];
    __shared__ uint32_t kept_total;

    const uint32_t c = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    constexpr uint32_t num_warps = kBoundedThreads / 32;

    const uint64_t g0 = args.group_offset[c];
    const uint32_t G = static_cast<uint32_t>(args.group_offset[c + 1] - g0);
    const uint64_t r0 = args.row_offset[c];
    const uint32_t rows = args.n_classes[c];
    const double * matrix = args.matrix + args.matrix_offset[c];
    const double * noise = args.noise + r0;
    const double * counts = args.counts + r0;
    const uint32_t * order = args.order + g0;
    const double * log_freq = args.log_freq + g0;

    const uint64_t p0 = args.pair_offset[c];
    const uint32_t n_pairs = static_cast<uint32_t>(args.pair_offset[c + 1] - p0);

    double * pair_ll = args.pair_ll + p0;
    double * raw = args.cand_loglik + p0;   // raw sums of all pairs first, compacted in place at the end

    // ---- values: a warp per pair
    for (uint32_t p = warp; p < n_pairs; p += num_warps) {

        uint32_t i, j;
        pairOfIndex(G, p, &i, &j);
        const uint32_t a = order[i], b = order[j];

        const double * col_a = matrix + static_cast<uint64_t>(a) * rows;
        const double * col_b = matrix + static_cast<uint64_t>(b) * rows;

        double acc = 0;
        for (uint32_t r = lane; r < rows; r += 32) {
            double prob = noise[r];
            prob += col_a[r] / 2.0;
            prob += col_b[r] / 2.0;
            acc += counts[r] * log(prob);
        }
        acc = warpSum(acc);

        if (lane == 0) {
            raw[p] = acc;
            pair_ll[p] = acc + (log_freq[a] + log_freq[b] + (a == b ? 0.0 : 0.6931471805599453));
        }
    }

    if (tid == 0) {
        carry = -DBL_MAX;
        kept_total = 0;
    }
    __syncthreads();

    // ---- exclusive prefix maximum in pair order, test, compaction (tiles of kBoundedThreads pairs)
    uint32_t * cand_pair = args.cand_pair + p0;

    for (uint32_t base = 0; base < n_pairs; base += kBoundedThreads) {

        const uint32_t p = base + tid;
        const double ll = p < n_pairs ? pair_ll[p] : -DBL_MAX;
        const double raw_value = p < n_pairs ? raw[p] : 0.0;

        // inclusive max within the warp, then across warps
        double incl = ll;
        #pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const double up = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) { incl = fmax(incl, up); }
        }
        double excl = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) { excl = -DBL_MAX; }

        if (lane == 31) { warp_max[warp] = incl; }
        __syncthreads();

        double before = carry;
        for (uint32_t w = 0; w < warp; ++w) { before = fmax(before, warp_max[w]); }
        excl = fmax(excl, before);

        const bool keep = p < n_pairs && !(ll - excl < args.min_log_diff);

        const uint32_t ballot = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) { warp_count[warp] = __popc(ballot); }
        __syncthreads();

        uint32_t offset = kept_total;
        for (uint32_t w = 0; w < warp; ++w) { offset += warp_count[w]; }
        offset += __popc(ballot & ((1u << lane) - 1));

        __syncthreads();   // all reads of raw[] / carry / kept_total of this tile are done

        if (keep) {
            cand_pair[offset] = p;
            raw[offset] = raw_value;   // offset <= p: never overwrites a value a later tile still needs
        }

        if (tid == kBoundedThreads - 1) {
            double tile_max = carry;
            for (uint32_t w = 0; w < num_warps; ++w) { tile_max = fmax(tile_max, warp_max[w]); }
            carry = tile_max;
            uint32_t total = kept_total;
            for (uint32_t w = 0; w < num_warps; ++w) { total += warp_count[w]; }
            kept_total = total;
        }
        __syncthreads();
    }

    if (tid == 0) { args.n_candidates[c] = kept_total; }
}

// candidates of every cluster, one after the other
__global__ void boundedGatherKernel(const uint64_t * __restrict__ pair_offset, const uint64_t * __restrict__ cand_offset, const uint32_t * __restrict__ cand_pair,
                                    const double * __restrict__ cand_loglik, uint32_t * __restrict__ out_pair, double * __restrict__ out_loglik) {

    const uint32_t c = blockIdx.x;
    const uint64_t src = pair_offset[c];
    const uint64_t dst = cand_offset[c];
    const uint32_t n = static_cast<uint32_t>(cand_offset[c + 1] - dst);

    for (uint32_t k = threadIdx.x; k < n; k += blockDim.x) {
        out_pair[dst + k] = cand_pair[src + k];
        out_loglik[dst + k] = cand_loglik[src + k];
    }
}

}

extern "C" int rpvg_b200_grouped_bounded_pairs(rpvg_b200_ctx * ctx, rpvg_b200_grouped * grouped, const uint64_t * group_offset, const uint32_t * order, const double * log_freq,
                                               double min_log_diff, const uint64_t ** cand_offset, const uint32_t ** cand_pair, const double ** cand_loglik) {

    if (!ctx || !grouped || !group_offset || !order || !log_freq || !cand_offset || !cand_pair || !cand_loglik) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }

    const uint32_t C = grouped->n_clusters;
    const uint64_t G = group_offset[C];

    std::vector<uint64_t> pair_offset(C + 1, 0);
    for (uint32_t c = 0; c < C; ++c) {
        const uint64_t g = group_offset[c + 1] - group_offset[c];
        if (g >= 65536) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_grouped_bounded_pairs: more than 65535 groups in a cluster"); }
        pair_offset[c + 1] = pair_offset[c] + g * (g + 1) / 2;
    }
    const uint64_t NP = pair_offset[C];

    grouped->h_cand_offset.assign(C + 1, 0);
    grouped->h_cand_pair.clear();
    grouped->h_cand_loglik.clear();
    *cand_offset = grouped->h_cand_offset.data();
    *cand_pair = nullptr;
    *cand_loglik = nullptr;

    if (NP == 0) { return RPVG_B200_OK; }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 2 * pad(8 * (C + 1)) + pad(4 * G) + pad(8 * G) + 2 * pad(8 * NP) + 2 * pad(4 * NP) + pad(4 * C) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocAsync(reinterpret_cast<void **>(&base), bytes, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_pair_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint32_t * d_order = reinterpret_cast<uint32_t *>(take(4 * G));
    double * d_log_freq = reinterpret_cast<double *>(take(8 * G));
    double * d_pair_ll = reinterpret_cast<double *>(take(8 * NP));
    double * d_cand_loglik = reinterpret_cast<double *>(take(8 * NP));
    uint32_t * d_cand_pair = reinterpret_cast<uint32_t *>(take(4 * NP));
    uint32_t * d_n_candidates = reinterpret_cast<uint32_t *>(take(4 * C));
    uint64_t * d_cand_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint32_t * d_out_pair = reinterpret_cast<uint32_t *>(take(4 * NP));
    double * d_out_loglik = d_pair_ll;   // free once the first kernel is through

    cudaError_t err = cudaMemcpyAsync(d_pair_offset, pair_offset.data(), 8 * (C + 1), cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) { err = cudaMemcpyAsync(d_order, order, 4 * G, cudaMemcpyHostToDevice, s); }
    if (err == cudaSuccess) { err = cudaMemcpyAsync(d_log_freq, log_freq, 8 * G, cudaMemcpyHostToDevice, s); }

    std::vector<uint32_t> n_candidates(C, 0);

    if (err == cudaSuccess) {

        BoundedArgs args;
        args.matrix = grouped->d_matrix;
        args.noise = grouped->d_noise;
        args.counts = grouped->d_counts;
        args.row_offset = grouped->d_row_offset;
        args.group_offset = grouped->d_group_offset;
        args.matrix_offset = grouped->d_matrix_offset;
        args.n_classes = grouped->d_n_classes;
        args.pair_offset = d_pair_offset;
        args.order = d_order;
        args.log_freq = d_log_freq;
        args.pair_ll = d_pair_ll;
        args.min_log_diff = min_log_diff;
        args.n_candidates = d_n_candidates;
        args.cand_pair = d_cand_pair;
        args.cand_loglik = d_cand_loglik;

        boundedPairsKernel<<<C, kBoundedThreads, 0, s>>>(args);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess) { err = cudaMemcpyAsync(n_candidates.data(), d_n_candidates, 4 * C, cudaMemcpyDeviceToHost, s); }
    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

    if (err == cudaSuccess) {

        for (uint32_t c = 0; c < C; ++c) { grouped->h_cand_offset[c + 1] = grouped->h_cand_offset[c] + n_candidates[c]; }

        grouped->h_cand_pair.resize(grouped->h_cand_offset[C]);
        grouped->h_cand_loglik.resize(grouped->h_cand_offset[C]);

        const uint64_t total = grouped->h_cand_offset[C];

        if (total > 0) {

            err = cudaMemcpyAsync(d_cand_offset, grouped->h_cand_offset.data(), 8 * (C + 1), cudaMemcpyHostToDevice, s);

            if (err == cudaSuccess) {
                boundedGatherKernel<<<C, 128, 0, s>>>(d_pair_offset, d_cand_offset, d_cand_pair, d_cand_loglik, d_out_pair, d_out_loglik);
                ctx->launches++;
                err = cudaGetLastError();
            }

            if (err == cudaSuccess) { err = cudaMemcpyAsync(grouped->h_cand_pair.data(), d_out_pair, 4 * total, cudaMemcpyDeviceToHost, s); }
            if (err == cudaSuccess) { err = cudaMemcpyAsync(grouped->h_cand_loglik.data(), d_out_loglik, 8 * total, cudaMemcpyDeviceToHost, s); }
        }

        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }
    }

    cudaFreeAsync(base, s);

    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_grouped_bounded_pairs: ") + cudaGetErrorString(err)); }

    *cand_pair = grouped->h_cand_pair.data();
    *cand_loglik = grouped->h_cand_loglik.data();
    return RPVG_B200_OK;
}
