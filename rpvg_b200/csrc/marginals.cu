// Haplotype probabilities and read counts per path from the estimates of the haplotype-transcripts model — the reducer
// the haplotype abundance table applies before printing.
//
// Reference semantics restated (never its code): HaplotypeAbundanceEstimatesWriter::addEstimates
// (rpvg:src/threaded_output_writer.cpp:355-391). A cluster's estimates are a list of path sets (each up to `ploidy`
// path indices, repeats adjacent), one posterior per set and one abundance per set member. Per path of the cluster:
//   read_count[p]       = sum of the abundances of every set member equal to p,
//   haplotype_prob[p]   = sum of the posteriors of the sets p is in (a repeated member counts once),
// both accumulated in the order of the sets.
//
// Device side, bit-identical to that order: one thread per set writes its members' keys (global path index) and the two
// contributions (a repeated member contributes posterior 0.0, an exact no-op on a non-negative sum); a stable radix sort
// of the member positions by key keeps the set order inside each path; one thread per path then adds its segment front
// to back. HBM-bound integer/byte work: ~40 B per member over the three passes plus the sort.

#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

__global__ void marginalKeysKernel(const uint32_t n_clusters, const uint64_t n_sets, const uint64_t * __restrict__ path_offset, const uint64_t * __restrict__ set_offset,
                                   const uint64_t * __restrict__ member_offset, const uint32_t * __restrict__ members, const double * __restrict__ posteriors,
                                   uint32_t * __restrict__ key, uint32_t * __restrict__ position, double * __restrict__ prob_part, uint32_t * __restrict__ error) {

    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n_sets) { return; }

    const uint32_t c = clusterOf(set_offset, n_clusters, i);
    const uint64_t p0 = path_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - p0;

    const double posterior = posteriors[i];
    uint32_t previous = 0xffffffffu;

    for (uint64_t e = member_offset[i]; e < member_offset[i + 1]; ++e) {

        const uint32_t member = members[e];
        if (member >= num_paths) {
            atomicOr(error, 1u);
            key[e] = static_cast<uint32_t>(p0);
            position[e] = static_cast<uint32_t>(e);
            prob_part[e] = 0.0;
            continue;
        }

        key[e] = static_cast<uint32_t>(p0 + member);
        position[e] = static_cast<uint32_t>(e);
        prob_part[e] = (member != previous) ? posterior : 0.0;
        previous = member;
    }
}

// segment of path p in the sorted keys = [lower_bound(p), lower_bound(p + 1))
__global__ void marginalSumKernel(const uint64_t n_paths, const uint64_t n_members, const uint32_t * __restrict__ sorted_key, const uint32_t * __restrict__ sorted_position,
                                  const double * __restrict__ prob_part, const double * __restrict__ abundances, double * __restrict__ haplotype_probs,
                                  double * __restrict__ read_counts) {

    const uint64_t p = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (p >= n_paths) { return; }

    uint64_t lo = 0, hi = n_members;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (sorted_key[mid] < p) { lo = mid + 1; } else { hi = mid; }
    }

    double prob = 0, count = 0;

    for (uint64_t k = lo; k < n_members && sorted_key[k] == p; ++k) {
        const uint32_t e = sorted_position[k];
        prob += prob_part[e];
        count += abundances[e];
    }

    haplotype_probs[p] = prob;
    read_counts[p] = count;
}

}

extern "C" int rpvg_b200_haplotype_marginals(rpvg_b200_ctx * ctx, uint32_t n_clusters, const uint64_t * path_offset, const uint64_t * set_offset, const uint64_t * member_offset,
                                             const uint32_t * members, const double * posteriors, const double * abundances, double * haplotype_probs, double * read_counts) {

    if (!ctx) { return RPVG_B200_ERR_INVALID; }
    if (n_clusters == 0) { return RPVG_B200_OK; }
    if (!path_offset || !set_offset || !member_offset || !haplotype_probs || !read_counts) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_haplotype_marginals: NULL argument"); }

    const uint64_t P = path_offset[n_clusters];
    const uint64_t S = set_offset[n_clusters];
    const uint64_t E = member_offset[S];

    if (path_offset[0] != 0 || set_offset[0] != 0 || member_offset[0] != 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_haplotype_marginals: offset arrays must start at 0"); }
    if (P >= 0xffffffffull || E >= 0x7fffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_haplotype_marginals: paths and set members must each stay below 2^31 per call"); }
    if (E > 0 && (!members || !posteriors || !abundances)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_haplotype_marginals: NULL array"); }

    if (P == 0) { return RPVG_B200_OK; }

    if (E == 0) {
        for (uint64_t p = 0; p < P; ++p) {   // nothing to add up: the reference prints zeros
            haplotype_probs[p] = 0;
            read_counts[p] = 0;
        }
        return RPVG_B200_OK;
    }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    int key_bits = 1;
    while ((uint64_t(1) << key_bits) < P && key_bits < 32) { key_bits++; }

    size_t sort_bytes = 0;
    RPVG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                                   static_cast<uint32_t *>(nullptr), static_cast<int>(E), 0, key_bits, s));

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = pad(8 * (n_clusters + 1)) * 2 + pad(8 * (S + 1)) + pad(4 * E) * 5 + pad(8 * S) + pad(8 * E) * 2 + pad(8 * P) * 2 + pad(sort_bytes) + 256;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));
    const DeviceAllocationGuard allocation_guard{s, base};

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (n_clusters + 1)));
    uint64_t * d_set_offset = reinterpret_cast<uint64_t *>(take(8 * (n_clusters + 1)));
    uint64_t * d_member_offset = reinterpret_cast<uint64_t *>(take(8 * (S + 1)));
    uint32_t * d_members = reinterpret_cast<uint32_t *>(take(4 * E));
    uint32_t * d_key = reinterpret_cast<uint32_t *>(take(4 * E));
    uint32_t * d_key_sorted = reinterpret_cast<uint32_t *>(take(4 * E));
    uint32_t * d_position = reinterpret_cast<uint32_t *>(take(4 * E));
    uint32_t * d_position_sorted = reinterpret_cast<uint32_t *>(take(4 * E));
    double * d_posteriors = reinterpret_cast<double *>(take(8 * S));
    double * d_abundances = reinterpret_cast<double *>(take(8 * E));
    double * d_prob_part = reinterpret_cast<double *>(take(8 * E));
    double * d_probs = reinterpret_cast<double *>(take(8 * P));
    double * d_counts = reinterpret_cast<double *>(take(8 * P));
    void * d_sort = take(sort_bytes);
    uint32_t * d_error = reinterpret_cast<uint32_t *>(take(4));

    cudaError_t err = cudaSuccess;
    #define RPVG_STEP(call) if (err == cudaSuccess) { err = (call); }

    RPVG_STEP(cudaMemcpyAsync(d_path_offset, path_offset, 8 * (n_clusters + 1), cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_set_offset, set_offset, 8 * (n_clusters + 1), cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_member_offset, member_offset, 8 * (S + 1), cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_members, members, 4 * E, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_posteriors, posteriors, 8 * S, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemcpyAsync(d_abundances, abundances, 8 * E, cudaMemcpyHostToDevice, s))
    RPVG_STEP(cudaMemsetAsync(d_error, 0, 4, s))

    if (err == cudaSuccess) {
        marginalKeysKernel<<<static_cast<unsigned>((S + 255) / 256), 256, 0, s>>>(n_clusters, S, d_path_offset, d_set_offset, d_member_offset, d_members, d_posteriors, d_key, d_position,
                                                                                d_prob_part, d_error);
        RPVG_LAUNCHED(ctx);
        err = cudaGetLastError();
    }

    RPVG_STEP(cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, d_key, d_key_sorted, d_position, d_position_sorted, static_cast<int>(E), 0, key_bits, s))
    if (err == cudaSuccess) { ctx->launches += 2 + static_cast<uint64_t>((key_bits + 7) / 8); }

    if (err == cudaSuccess) {
        marginalSumKernel<<<static_cast<unsigned>((P + 255) / 256), 256, 0, s>>>(P, E, d_key_sorted, d_position_sorted, d_prob_part, d_abundances, d_probs, d_counts);
        RPVG_LAUNCHED(ctx);
        err = cudaGetLastError();
    }

    uint32_t dev_error = 0;
    RPVG_STEP(cudaMemcpyAsync(haplotype_probs, d_probs, 8 * P, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaMemcpyAsync(read_counts, d_counts, 8 * P, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaMemcpyAsync(&dev_error, d_error, 4, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaStreamSynchronize(s))
    #undef RPVG_STEP


    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_haplotype_marginals: ") + cudaGetErrorString(err)); }
    if (dev_error) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_haplotype_marginals: set member out of the cluster's path range"); }

    return RPVG_B200_OK;
}
