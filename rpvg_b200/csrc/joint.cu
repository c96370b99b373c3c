// Read counts and TPM per member of every haplotype combination — the reducer behind the joint haplotype abundance table.
//
// Reference semantics restated (never its code): JointHaplotypeAbundanceEstimatesWriter::addEstimates
// (rpvg:src/threaded_output_writer.cpp:456-525). A cluster's estimates are a list of path sets with one posterior per
// set and a flat list of abundances. Only sets with posterior >= min_posterior are written, and the reference's
// abundance iterator advances only over the members of WRITTEN sets (:473-502): the k-th abundance of the cluster belongs
// to the k-th member of the written sets, not to the k-th member overall. Per written member:
//   read_count = that abundance, tpm = (abundance / effective_length, 0 without effective length) / total_transcript_count * 10^6.
// noise_counts: every one of the `ploidy` slots receives noise_count_c / ploidy, cluster after cluster (:517-520).
//
// Device side: one warp per cluster walks its sets 32 at a time (exclusive warp scan of the written sets' sizes = the
// iterator's position), lanes then write their own members; the noise sum is one thread adding in cluster order, so that
// every value is bit-identical to the reference's. HBM-bound: ~30 B per member.

#include "common.cuh"
#include "em_batch.cuh"

namespace {

__global__ void jointEstimatesKernel(const uint32_t n_clusters, const uint64_t * __restrict__ path_offset, const uint64_t * __restrict__ set_offset,
                                     const uint64_t * __restrict__ member_offset, const uint32_t * __restrict__ members, const double * __restrict__ posteriors,
                                     const double * __restrict__ abundances, const double * __restrict__ effective_lengths, const double min_posterior,
                                     const double total_transcript_count, uint8_t * __restrict__ written, double * __restrict__ read_counts, double * __restrict__ tpm,
                                     uint32_t * __restrict__ error) {

    const uint32_t lane = threadIdx.x & 31;
    const uint64_t c = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
    if (c >= n_clusters) { return; }

    const uint64_t s0 = set_offset[c], s1 = set_offset[c + 1];
    const uint64_t p0 = path_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - p0;
    const uint64_t block = member_offset[s0];   // the cluster's abundances start where its members do

    uint64_t consumed = 0;   // abundances the reference's iterator has passed

    for (uint64_t base = s0; base < s1; base += 32) {

        const uint64_t i = base + lane;
        const bool live = i < s1;

        const uint64_t e0 = live ? member_offset[i] : 0;
        const uint32_t size = live ? static_cast<uint32_t>(member_offset[i + 1] - e0) : 0;
        const bool keep = live && posteriors[i] >= min_posterior;

        uint32_t incl = keep ? size : 0;
        #pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) { incl += up; }
        }

        const uint64_t at = block + consumed + incl - (keep ? size : 0);
        consumed += __shfl_sync(0xffffffffu, incl, 31);

        if (!live) { continue; }

        written[i] = keep ? 1 : 0;

        for (uint32_t j = 0; j < size; ++j) {

            double count = 0, per_million = 0;

            if (keep) {

                const uint32_t member = members[e0 + j];
                if (member >= num_paths) {
                    atomicOr(error, 1u);
                    continue;
                }

                count = abundances[at + j];
                const double length = effective_lengths[p0 + member];
                const double transcript_count = length > 0 ? count / length : 0.0;
                per_million = transcript_count / total_transcript_count * 1e6;
            }

            read_counts[e0 + j] = count;
            tpm[e0 + j] = per_million;
        }
    }
}

__global__ void jointNoiseKernel(const uint32_t n_clusters, const double * __restrict__ noise_count, const uint32_t ploidy, double * __restrict__ noise_counts) {

    if (blockIdx.x != 0 || threadIdx.x != 0) { return; }

    double sum = 0;
    for (uint32_t c = 0; c < n_clusters; ++c) { sum += noise_count[c] / ploidy; }
    for (uint32_t k = 0; k < ploidy; ++k) { noise_counts[k] = sum; }
}

}

extern "C" int rpvg_b200_joint_haplotype_estimates(rpvg_b200_ctx * ctx, uint32_t n_clusters, const uint64_t * path_offset, const uint64_t * set_offset,
                                                   const uint64_t * member_offset, const uint32_t * members, const double * posteriors, const double * abundances,
                                                   const double * effective_lengths, const double * noise_count, uint32_t ploidy, double min_posterior,
                                                   double total_transcript_count, uint8_t * written, double * read_counts, double * tpm, double * noise_counts) {

    if (!ctx) { return RPVG_B200_ERR_INVALID; }
    if (ploidy == 0 || ploidy > 64) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_joint_haplotype_estimates: ploidy must be 1..64"); }

    if (n_clusters == 0) {
        for (uint32_t k = 0; noise_counts && k < ploidy; ++k) { noise_counts[k] = 0; }
        return RPVG_B200_OK;
    }

    if (!path_offset || !set_offset || !member_offset || !noise_count || !noise_counts) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_joint_haplotype_estimates: NULL argument"); }
    if (path_offset[0] != 0 || set_offset[0] != 0 || member_offset[0] != 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_joint_haplotype_estimates: offset arrays must start at 0"); }

    const uint64_t P = path_offset[n_clusters];
    const uint64_t S = set_offset[n_clusters];
    const uint64_t E = member_offset[S];

    if (S > 0 && (!posteriors || !written)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_joint_haplotype_estimates: NULL array"); }
    if (E > 0 && (!members || !abundances || !effective_lengths || !read_counts || !tpm)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_joint_haplotype_estimates: NULL array"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    const size_t bytes = 2 * padded<uint64_t>(n_clusters + 1) + padded<uint64_t>(S + 1) + padded<uint32_t>(E) + padded<double>(S) + 3 * padded<double>(E) + padded<double>(P) +
                         padded<double>(n_clusters) + padded<uint8_t>(S) + padded<double>(ploidy) + padded<uint32_t>(1) + 4096;

    DevicePool pool;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&pool.base), bytes, ctx->pool, s));
    const DeviceAllocationGuard allocation_guard{s, pool.base};
    pool.size = bytes;

    uint64_t * d_path_offset = pool.take<uint64_t>(n_clusters + 1);
    uint64_t * d_set_offset = pool.take<uint64_t>(n_clusters + 1);
    uint64_t * d_member_offset = pool.take<uint64_t>(S + 1);
    uint32_t * d_members = pool.take<uint32_t>(E);
    double * d_posteriors = pool.take<double>(S);
    double * d_abundances = pool.take<double>(E);
    double * d_counts = pool.take<double>(E);
    double * d_tpm = pool.take<double>(E);
    double * d_lengths = pool.take<double>(P);
    double * d_noise = pool.take<double>(n_clusters);
    uint8_t * d_written = pool.take<uint8_t>(S);
    double * d_noise_counts = pool.take<double>(ploidy);
    uint32_t * d_error = pool.take<uint32_t>(1);

    cudaError_t err = cudaSuccess;
    auto up = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); }
    };

    up(d_path_offset, path_offset, 8 * (n_clusters + 1));
    up(d_set_offset, set_offset, 8 * (n_clusters + 1));
    up(d_member_offset, member_offset, 8 * (S + 1));
    up(d_members, members, 4 * E);
    up(d_posteriors, posteriors, 8 * S);
    up(d_abundances, abundances, 8 * E);
    up(d_lengths, effective_lengths, 8 * P);
    up(d_noise, noise_count, 8 * static_cast<size_t>(n_clusters));
    if (err == cudaSuccess) { err = cudaMemsetAsync(d_error, 0, 4, s); }

    if (err == cudaSuccess) {

        if (S > 0) {
            const unsigned grid = static_cast<unsigned>((static_cast<uint64_t>(n_clusters) * 32 + 255) / 256);
            jointEstimatesKernel<<<grid, 256, 0, s>>>(n_clusters, d_path_offset, d_set_offset, d_member_offset, d_members, d_posteriors, d_abundances, d_lengths, min_posterior,
                                                      total_transcript_count, d_written, d_counts, d_tpm, d_error);
            RPVG_LAUNCHED(ctx);
        }

        jointNoiseKernel<<<1, 32, 0, s>>>(n_clusters, d_noise, ploidy, d_noise_counts);
        RPVG_LAUNCHED(ctx);
        err = cudaGetLastError();
    }

    uint32_t dev_error = 0;
    auto down = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, s); }
    };

    down(written, d_written, S);
    down(read_counts, d_counts, 8 * E);
    down(tpm, d_tpm, 8 * E);
    down(noise_counts, d_noise_counts, 8 * static_cast<size_t>(ploidy));
    down(&dev_error, d_error, 4);
    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }


    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_joint_haplotype_estimates: ") + cudaGetErrorString(err)); }
    if (dev_error) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_joint_haplotype_estimates: set member out of the cluster's path range"); }

    return RPVG_B200_OK;
}
