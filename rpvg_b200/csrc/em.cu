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
//   scratch                                         per-CTA work area of the EM kernel (em_sell.cuh)
// The reference is dense and column-major; here the matrix is sparse in both orientations so that the E-step (row sums)
// and the M-step (column sums) are both gathers with a fixed summation order: no atomics, results are run-to-run
// deterministic, which matters because the stopping rule is a 1e-3 step test while parity is 1e-6.

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "em_kernels.cuh"
#include "em_sell.cuh"
#include "em_dense.cuh"
#include "em_stream.cuh"
#include "em_batch.cuh"

namespace {

// ---- build ----------------------------------------------------------------------------------------------------------
// The read classes of a cluster are re-ordered by their first (lowest) path index: EM does not care about the order of
// read classes, and with this order a contiguous slice of them touches a narrow band of paths, which keeps the
// cross-CTA exchange of the EM kernel short (em_sell.cuh). Sort key: (cluster, first path); empty read classes last.
__global__ void rowKeysKernel(uint32_t n_clusters, uint64_t n_reads, const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ row_ptr,
                              const uint32_t * __restrict__ col, uint64_t * __restrict__ key, uint32_t * __restrict__ row_id) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_reads) { return; }

    const uint32_t c = clusterOf(read_offset, n_clusters, r);
    const uint32_t first = (row_ptr[r + 1] > row_ptr[r]) ? col[row_ptr[r]] : 0xffffffffu;

    key[r] = (static_cast<uint64_t>(c) << 32) | first;
    row_id[r] = static_cast<uint32_t>(r);
}

// Length of every read class in sorted order (one extra zero at the end so that the scan yields n_reads + 1 offsets).
__global__ void sortedLengthsKernel(uint64_t n_reads, const uint32_t * __restrict__ perm, const uint64_t * __restrict__ row_ptr, uint64_t * __restrict__ len) {

    const uint64_t p = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (p > n_reads) { return; }

    uint64_t l = 0;
    if (p < n_reads) {
        const uint32_t r = perm[p];
        l = row_ptr[r + 1] - row_ptr[r];
    }
    len[p] = l;
}

// One thread per read class (sorted position p, original row perm[p]).
// Row normalisation exactly as the reference orders it: v = p / rowsum (left-to-right sum over ascending path index,
// which is what the dense row-wise sum does since zeros add nothing), then v * (1 - noise), NaN -> 0.
__global__ void buildRowsKernel(uint32_t n_clusters, uint64_t n_reads,
                                const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ path_offset,
                                const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col,
                                const double * __restrict__ prob, const double * __restrict__ noise, const uint32_t * __restrict__ count,
                                const uint32_t * __restrict__ perm, const uint64_t * __restrict__ sorted_row_ptr,
                                uint32_t * __restrict__ sorted_col, double * __restrict__ sorted_noise,
                                double * __restrict__ val, double * __restrict__ cnt, uint32_t * __restrict__ entry_row,
                                uint32_t * __restrict__ key, uint32_t * __restrict__ idx, uint32_t * __restrict__ error,
                                const double * __restrict__ count_f64, const int prenormalised) {

    const uint64_t p = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (p >= n_reads) { return; }

    const uint64_t r = perm[p];

    // sorting keeps clusters contiguous, so position p lies in the same cluster's range as row r
    const uint32_t c = clusterOf(read_offset, n_clusters, p);
    const uint64_t path_base = path_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - path_base;
    const uint32_t local_row = static_cast<uint32_t>(p - read_offset[c]);

    const uint64_t e0 = row_ptr[r];
    const uint64_t e1 = row_ptr[r + 1];
    const uint64_t out = sorted_row_ptr[p];

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

        double v = prob[e];
        if (!prenormalised) {
            v = (v / row_sum) * scale;
            if (isnan(v)) { v = 0; }
        }

        const uint64_t o = out + (e - e0);
        const uint32_t j = static_cast<uint32_t>(min(static_cast<uint64_t>(col[e]), num_paths - 1));

        val[o] = v;
        sorted_col[o] = j;
        entry_row[o] = local_row;
        key[o] = static_cast<uint32_t>(path_base + j);
        idx[o] = static_cast<uint32_t>(o);
    }

    sorted_noise[p] = noise[r];
    cnt[p] = prenormalised ? count_f64[r] : static_cast<double>(count[r]);

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

// Holds a stream back for a fixed time (one thread, no memory traffic): see the launch loop of rpvg_b200_batch_em.
__global__ void spacerKernel(unsigned long long ns) {

    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    do {
        __nanosleep(1000);
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    } while (t1 - t0 < ns);
}

uint32_t envKnob(const char * name, uint32_t fallback) {

    const char * v = std::getenv(name);
    return (v && *v) ? static_cast<uint32_t>(std::strtoul(v, nullptr, 10)) : fallback;
}

constexpr int kNumTiers = 5;

// sides of the dense kernels (em_dense.cuh), descending; a cluster goes to the smallest that holds max(M, N + 1)
constexpr uint32_t kDenseDims[4] = {128, 64, 32, 16};

// Largest side that goes to the dense kernels unless RPVG_B200_EM_DENSE_MAX says otherwise. Measured on B200 (config 2,
// per-CTA timeline): 0.33-0.39 us per iteration at side 16 and 0.52-0.55 at 32, against 1.1-1.3 on the sliced-ELL
// kernels; at 64 (1.15-1.4 vs 1.6-1.9) the gain no longer pays for the shared memory and at 128 there is none.
constexpr uint32_t kDenseDefaultMax = 32;

// Shared-memory tiers of the EM launches and the CTA size that goes with each; RPVG_B200_EM_TIERS="kb:threads,..." (five
// pairs, ascending) overrides them for tuning runs.
struct SellTiers {

    // (two CTAs of the 104 KB tier, four of the 52 KB tier per SM; with ~125 registers per thread an SM holds 512 threads)
    uint32_t smem[kNumTiers] = {8 * 1024, 24 * 1024, 52 * 1024, 104 * 1024, 216 * 1024};
    uint32_t threads[kNumTiers] = {64, 128, 128, 256, 512};

    SellTiers() {
        const char * v = std::getenv("RPVG_B200_EM_TIERS");
        if (!v || !*v) { return; }
        unsigned kb[kNumTiers], th[kNumTiers];
        if (std::sscanf(v, "%u:%u,%u:%u,%u:%u,%u:%u,%u:%u", &kb[0], &th[0], &kb[1], &th[1], &kb[2], &th[2], &kb[3], &th[3], &kb[4], &th[4]) == 2 * kNumTiers) {
            for (int t = 0; t < kNumTiers; ++t) { smem[t] = kb[t] * 1024; threads[t] = th[t]; }
        }
    }
};

const SellTiers & sellTiers() {
    static const SellTiers tiers;
    return tiers;
}
constexpr uint32_t kStaticSmemReserve = 12 * 1024;   // static shared memory of emSellBuildKernel<512> + slack

struct GroupPlan {

    uint32_t cluster;
    uint64_t weight;
    int bin;              // sort key: group size (descending), tier (descending), index width
    uint64_t cta_bytes;   // scratch bytes per CTA
};

// Bins the clusters of a batch into launches. Within a bin, largest first, like the reference's schedule
// (rpvg:src/main.cpp:827): the hardware block scheduler then hands the long-running clusters out first.
void planEmLaunches(const rpvg_b200_ctx * ctx, rpvg_b200_batch * b, std::vector<uint32_t> * order, std::vector<uint64_t> * group_scratch,
                    std::vector<uint64_t> * group_cta_bytes) {

    // weight (entries + rows) one CTA should own; above it a cluster is split over 2, 4, 8, 16 CTAs
    const uint32_t slice_weight = std::max(64u, envKnob("RPVG_B200_EM_SLICE_WEIGHT", 12000));
    const uint32_t big_threads = envKnob("RPVG_B200_EM_BIG_THREADS", 1024);
    const uint32_t max_group = std::min(ctx->max_group_size, std::max(1u, envKnob("RPVG_B200_EM_MAX_GROUP", kMaxGroupSize)));
    const uint32_t smem_cap = std::min<uint32_t>(sellTiers().smem[kNumTiers - 1], static_cast<uint32_t>(ctx->max_smem_optin) - kStaticSmemReserve);

    std::vector<GroupPlan> items(b->n_clusters);

    // bin index = ((4 - log2 group) * tiers + (tiers - 1 - tier)) * 2 + (index_bytes == 4); then the fallback, then the
    // dense kernels (largest first)
    const int fallback_bin = 5 * kNumTiers * 2;
    const int dense_bin0 = fallback_bin + 1;
    const int stream_bin = -1;   // first: one cooperative launch per cluster, before anything else is forked

    // Above this weight a cluster gets the whole GPU (emStreamKernel): 16 slices of shared memory no longer hold it and
    // what does not fit would be streamed through 16 SMs only. Also for clusters whose vectors alone outgrow a CTA.
    const uint64_t stream_weight = std::max<uint64_t>(1000, envKnob("RPVG_B200_EM_STREAM_WEIGHT", 400000));
    const uint32_t dense_max = std::min(envKnob("RPVG_B200_EM_DENSE_MAX", kDenseDefaultMax), kDenseDims[0]);

    for (uint32_t c = 0; c < b->n_clusters; ++c) {

        const uint64_t M = b->h_read_offset[c + 1] - b->h_read_offset[c];
        const uint64_t N = b->h_path_offset[c + 1] - b->h_path_offset[c];
        const uint64_t Zc = b->h_cluster_nnz[c];
        const uint64_t weight = Zc + M;

        uint32_t group = 1;
        while (group < max_group && weight > static_cast<uint64_t>(slice_weight) * group) { group *= 2; }

        // bounds of one slice (slices are cut at equal weight)
        const uint64_t m_b = std::min(M, weight / group + 2);
        const uint64_t z_b = std::min(Zc, weight / group + std::min(N, Zc) + 2);
        const uint64_t index_bytes = (N + 1 < 65536 && m_b < 65536) ? 2 : 4;

        const bool multi = group > 1;

        // columns a slice has entries in: all of them at worst; with the read classes ordered by first path a slice
        // of a group touches a band of about N / group columns
        const uint64_t cols_worst = std::min(N, z_b);
        const uint64_t cols_typical = multi ? std::min(cols_worst, 2 * N / group + 64) : cols_worst;

        // what must be in shared memory whatever the layout: the abundances (one CTA) or the two generations of column
        // sums the peers read (group); a group also needs N bytes of exchange area while it is being set up
        const uint64_t fixed = multi ? std::max<uint64_t>(2 * sellRound16(8 * (cols_worst + 2)), N) : sellRound16(8 * (N + 1));

        GroupPlan plan;
        plan.cluster = c;
        plan.weight = weight;
        plan.bin = fallback_bin;
        plan.cta_bytes = 0;

        // small clusters: dense in shared memory, one thread per read class / path (em_dense.cuh)
        const uint64_t dense_side = std::max(M, N + 1);

        if (ctx->stream_grid > 0 && (weight > stream_weight || (fixed + 4096 > smem_cap && weight > 50000))) {

            plan.bin = stream_bin;

        } else if (dense_side <= dense_max) {

            int k = 3;
            while (kDenseDims[k] < dense_side) { k--; }
            plan.bin = dense_bin0 + k;

        } else if (fixed + 4096 <= smem_cap) {

            const uint64_t need_typical = std::max(fixed + 4096, sellSmemBytes(m_b, z_b, N, cols_typical, index_bytes, multi, false));
            const uint64_t need_worst = std::max(fixed + 4096, sellSmemBytes(m_b, z_b, N, cols_worst, index_bytes, multi, true));

            int tier_typical = kNumTiers - 1;
            for (int t = 0; t < kNumTiers; ++t) {
                if (need_typical <= std::min(sellTiers().smem[t], smem_cap)) { tier_typical = t; break; }
            }

            int tier_worst = kNumTiers;
            for (int t = 0; t < kNumTiers; ++t) {
                if (need_worst <= std::min(sellTiers().smem[t], smem_cap)) { tier_worst = t; break; }
            }

            // one tier up buys the guarantee that nothing spills; beyond that the slice is given scratch for its entries
            int tier = tier_typical;
            bool with_entries = true;
            if (tier_worst < kNumTiers && tier_worst <= tier_typical + 1) {
                tier = tier_worst;
                with_entries = false;
            }

            int group_log2 = 0;
            while ((1u << group_log2) < group) { group_log2++; }

            plan.bin = ((4 - group_log2) * kNumTiers + (kNumTiers - 1 - tier)) * 2 + (index_bytes == 4 ? 1 : 0);
            plan.cta_bytes = sellScratchBytes(m_b, z_b, N, index_bytes, with_entries) + sellHandoverBytes(std::min(sellTiers().smem[tier], smem_cap));
        }

        items[c] = plan;
    }

    std::stable_sort(items.begin(), items.end(), [](const GroupPlan & x, const GroupPlan & y) {
        if (x.bin != y.bin) { return x.bin < y.bin; }
        return x.weight > y.weight;
    });

    order->resize(b->n_clusters);
    group_scratch->resize(b->n_clusters);
    group_cta_bytes->resize(b->n_clusters);
    b->bins.clear();
    b->scratch_bytes = 0;

    for (uint32_t k = 0; k < items.size(); ++k) {

        (*order)[k] = items[k].cluster;

        if (k == 0 || items[k].bin != items[k - 1].bin) {

            rpvg_b200_batch::EmBin bin;
            bin.first = k;
            bin.count = 0;

            bin.dense_dim = 0;
            bin.stream = false;

            if (items[k].bin == stream_bin) {

                bin.stream = true;
                bin.group_size = 0;
                bin.threads = kStreamThreads;
                bin.smem = 0;
                bin.index_bytes = 4;

            } else if (items[k].bin >= dense_bin0) {

                bin.dense_dim = kDenseDims[items[k].bin - dense_bin0];
                bin.group_size = 1;
                bin.threads = std::max(32u, bin.dense_dim);
                bin.smem = static_cast<uint32_t>(sizeof(double) * (bin.dense_dim * (bin.dense_dim + 1) + 2 * bin.dense_dim));
                bin.index_bytes = 0;

            } else if (items[k].bin == fallback_bin) {

                bin.group_size = 0;
                bin.threads = kEmThreads;
                bin.smem = 0;
                bin.index_bytes = 4;

            } else {

                const int code = items[k].bin / 2;
                const int tier = kNumTiers - 1 - (code % kNumTiers);

                bin.group_size = 1u << (4 - code / kNumTiers);
                bin.index_bytes = (items[k].bin & 1) ? 4 : 2;
                bin.smem = std::min(sellTiers().smem[tier], smem_cap);
                bin.threads = sellTiers().threads[tier];

                // instantiations that exist (build threads): one CTA {64,128,256,512} x u16; groups {256,512} x u16; 512 x u32
                if (tier == kNumTiers - 1 && big_threads == 1024) { bin.threads = 1024; }
                if (bin.index_bytes == 4) { bin.threads = 512; }
                if (bin.group_size > 1) { bin.threads = std::max(bin.threads, 256u); }
            }

            b->bins.push_back(bin);
        }

        b->bins.back().count++;

        const uint64_t group_bytes = items[k].cta_bytes * std::max(1u, b->bins.back().group_size);
        (*group_scratch)[k] = b->scratch_bytes;
        (*group_cta_bytes)[k] = items[k].cta_bytes;
        b->scratch_bytes += group_bytes;
    }
}

// One bin = two launches on the same stream with the same grid and thread-block clusters: the build kernel lays the
// slices out, the run kernel iterates on them (em_sell.cuh).
template <int BUILD_THREADS, int RUN_THREADS, bool MULTI, typename IndexT>
cudaError_t launchSell(const SellArgs & sa, uint32_t group_size, uint32_t n_groups, cudaStream_t stream) {

    cudaLaunchConfig_t config = {};
    config.gridDim = dim3(n_groups * group_size, 1, 1);
    config.blockDim = dim3(BUILD_THREADS, 1, 1);
    config.dynamicSmemBytes = sa.smem_bytes;
    config.stream = stream;

    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = group_size;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    config.attrs = attr;
    config.numAttrs = MULTI ? 1 : 0;

    SellArgs args = sa;
    args.run_warps = RUN_THREADS / 32;

    cudaError_t err = cudaLaunchKernelEx(&config, emSellBuildKernel<BUILD_THREADS, MULTI, IndexT>, args);
    if (err != cudaSuccess) { return err; }

    config.blockDim = dim3(RUN_THREADS, 1, 1);
    return cudaLaunchKernelEx(&config, emSellRunKernel<RUN_THREADS, MULTI, IndexT>, args);
}

cudaError_t launchSellBin(const SellArgs & sa, const rpvg_b200_batch::EmBin & bin, cudaStream_t stream) {

    const bool multi = bin.group_size > 1;

    if (bin.index_bytes == 4) {
        if (bin.threads == 1024) {
            return multi ? launchSell<512, 1024, true, uint32_t>(sa, bin.group_size, bin.count, stream) : launchSell<512, 1024, false, uint32_t>(sa, 1, bin.count, stream);
        }
        return multi ? launchSell<512, 512, true, uint32_t>(sa, bin.group_size, bin.count, stream) : launchSell<512, 512, false, uint32_t>(sa, 1, bin.count, stream);
    }

    if (multi) {
        if (bin.threads == 256) { return launchSell<256, 256, true, uint16_t>(sa, bin.group_size, bin.count, stream); }
        if (bin.threads == 1024) { return launchSell<512, 1024, true, uint16_t>(sa, bin.group_size, bin.count, stream); }
        return launchSell<512, 512, true, uint16_t>(sa, bin.group_size, bin.count, stream);
    }

    // Where shared memory, not threads, limits the CTAs of a tier per SM (4 x 52 KB, 2 x 104 KB) the run kernel gets
    // twice the build kernel's threads: 32 warps per SM again, and half the latency per iteration for a cluster that
    // is alone on its SM at the end of a batch (measured: 57.7 -> 57.0 ms/step on config 2).
    switch (bin.threads) {
        case 64: return launchSell<64, 64, false, uint16_t>(sa, 1, bin.count, stream);
        case 128:
            if (bin.smem > 32 * 1024) { return launchSell<128, 256, false, uint16_t>(sa, 1, bin.count, stream); }
            return launchSell<128, 128, false, uint16_t>(sa, 1, bin.count, stream);
        case 256: return launchSell<256, 512, false, uint16_t>(sa, 1, bin.count, stream);
        case 1024: return launchSell<512, 1024, false, uint16_t>(sa, 1, bin.count, stream);
        default: return launchSell<512, 512, false, uint16_t>(sa, 1, bin.count, stream);
    }
}

cudaError_t launchDenseBin(const EmArgs & args, const uint32_t * clusters, const rpvg_b200_batch::EmBin & bin, cudaStream_t stream) {

    switch (bin.dense_dim) {
        case 16: emDenseKernel<16><<<bin.count, DenseShape<16>::kThreads, DenseShape<16>::kSmemBytes, stream>>>(args, clusters); break;
        case 32: emDenseKernel<32><<<bin.count, DenseShape<32>::kThreads, DenseShape<32>::kSmemBytes, stream>>>(args, clusters); break;
        case 64: emDenseKernel<64><<<bin.count, DenseShape<64>::kThreads, DenseShape<64>::kSmemBytes, stream>>>(args, clusters); break;
        default: emDenseKernel<128><<<bin.count, DenseShape<128>::kThreads, DenseShape<128>::kSmemBytes, stream>>>(args, clusters); break;
    }
    return cudaGetLastError();
}

template <typename Kernel>
cudaError_t configureSellKernel(Kernel kernel, int threads, bool multi, int smem_cap, uint32_t * max_group) {

    cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap);
    if (err != cudaSuccess || !multi) { return err; }

    err = cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (err != cudaSuccess) { return err; }

    // largest thread-block cluster that can be co-scheduled with the biggest shared-memory tier
    uint32_t group = kMaxGroupSize;

    while (group > 1) {

        cudaLaunchConfig_t config = {};
        config.gridDim = dim3(group, 1, 1);
        config.blockDim = dim3(threads, 1, 1);
        config.dynamicSmemBytes = smem_cap;

        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = group;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        config.attrs = attr;
        config.numAttrs = 1;

        int num_clusters = 0;
        err = cudaOccupancyMaxActiveClusters(&num_clusters, kernel, &config);
        if (err == cudaSuccess && num_clusters > 0) { break; }

        cudaGetLastError();
        group /= 2;
    }

    *max_group = std::min(*max_group, group);
    return cudaSuccess;
}

template <int BUILD_THREADS, int RUN_THREADS, bool MULTI, typename IndexT>
cudaError_t configureSell(const rpvg_b200_ctx * ctx, uint32_t * max_group) {

    const int smem_cap = std::min<int>(sellTiers().smem[kNumTiers - 1], ctx->max_smem_optin - static_cast<int>(kStaticSmemReserve));

    cudaError_t err = configureSellKernel(emSellBuildKernel<BUILD_THREADS, MULTI, IndexT>, BUILD_THREADS, MULTI, smem_cap, max_group);
    if (err != cudaSuccess) { return err; }
    return configureSellKernel(emSellRunKernel<RUN_THREADS, MULTI, IndexT>, RUN_THREADS, MULTI, smem_cap, max_group);
}

}

// Shape checks every entry point that takes a cluster arena makes on the host before anything goes to the device.
int rpvgCheckClusters(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl) {

    if (!cl) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "clusters is NULL"); }
    if (cl->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "empty batch (n_clusters == 0)"); }
    if (!cl->read_offset || !cl->path_offset) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "offset arrays are NULL"); }
    if (cl->read_offset[cl->n_clusters] > 0 && !cl->row_ptr) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "row_ptr is NULL"); }
    if (cl->read_offset[0] != 0 || cl->path_offset[0] != 0 || (cl->row_ptr && cl->row_ptr[0] != 0)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "offset arrays must start at 0"); }

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
    const uint64_t nnz = n_reads > 0 ? cl->row_ptr[n_reads] : 0;

    // (the sorts and scans of the matrix build take 32-bit signed item counts)
    if (n_paths + cl->n_clusters >= 0x7fffffffull || nnz >= 0x7fffffffull || n_reads >= 0x7fffffffull) {
        return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "batch too large: paths, read classes and non-zeros must each stay below 2^31 per batch");
    }

    // entry offsets non-decreasing over all read classes: the device lays its arrays out from them
    for (uint64_t r = 0; r < n_reads; ++r) {
        if (cl->row_ptr[r + 1] < cl->row_ptr[r]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "malformed cluster arena: row_ptr decreases at read class " + std::to_string(r)); }
    }
    if (n_reads > 0 && (!cl->noise || !cl->count)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "noise/count arrays are NULL"); }
    if (nnz > 0 && (!cl->col || !cl->prob)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "col/prob arrays are NULL"); }

    return RPVG_B200_OK;
}

namespace {

}

int rpvgEmInit(rpvg_b200_ctx * ctx) {

    uint32_t max_group = kMaxGroupSize;
    uint32_t unused = kMaxGroupSize;

    RPVG_CUDA(ctx, (configureSell<64, 64, false, uint16_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<128, 128, false, uint16_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<128, 256, false, uint16_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<256, 512, false, uint16_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<512, 512, false, uint16_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<512, 1024, false, uint16_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<512, 512, false, uint32_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<512, 1024, false, uint32_t>(ctx, &unused)));
    RPVG_CUDA(ctx, (configureSell<256, 256, true, uint16_t>(ctx, &max_group)));
    RPVG_CUDA(ctx, (configureSell<512, 512, true, uint16_t>(ctx, &max_group)));
    RPVG_CUDA(ctx, (configureSell<512, 1024, true, uint16_t>(ctx, &max_group)));
    RPVG_CUDA(ctx, (configureSell<512, 512, true, uint32_t>(ctx, &max_group)));
    RPVG_CUDA(ctx, (configureSell<512, 1024, true, uint32_t>(ctx, &max_group)));

    RPVG_CUDA(ctx, cudaFuncSetAttribute(emDenseKernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(DenseShape<128>::kSmemBytes)));

    ctx->max_group_size = max_group;

    // cooperative grid of emStreamKernel: every CTA resident at once
    {
        int cooperative = 0, per_sm = 0;
        cudaDeviceGetAttribute(&cooperative, cudaDevAttrCooperativeLaunch, ctx->device);
        if (cooperative && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, emStreamKernel, kStreamThreads, 0) == cudaSuccess && per_sm > 0) {
            ctx->stream_grid = std::min<uint32_t>(kStreamMaxCtas, static_cast<uint32_t>(std::min(per_sm, 4) * ctx->num_sms));
        }
        cudaGetLastError();
    }

    return RPVG_B200_OK;
}

int rpvgBatchCreate(rpvg_b200_ctx * ctx, uint32_t n_clusters, const uint64_t * read_offset, const uint64_t * path_offset, const uint64_t * cluster_nnz,
                    bool prenormalised, rpvg_b200_batch ** out_batch) {

    *out_batch = nullptr;

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));

    auto * b = new rpvg_b200_batch();
    b->n_clusters = n_clusters;
    b->n_reads = read_offset[n_clusters];
    b->n_paths = path_offset[n_clusters];
    b->prenormalised = prenormalised;

    b->h_read_offset.assign(read_offset, read_offset + n_clusters + 1);
    b->h_path_offset.assign(path_offset, path_offset + n_clusters + 1);
    b->h_cluster_nnz.assign(cluster_nnz, cluster_nnz + n_clusters);

    b->nnz = 0;
    for (uint32_t c = 0; c < n_clusters; ++c) { b->nnz += cluster_nnz[c]; }

    std::vector<uint32_t> order;
    std::vector<uint64_t> group_scratch;
    std::vector<uint64_t> group_cta_bytes;
    planEmLaunches(ctx, b, &order, &group_scratch, &group_cta_bytes);
    b->h_order = order;

    const uint64_t C = b->n_clusters, R = b->n_reads, P = b->n_paths, Z = b->nnz;

    b->sort_bits = 1;
    while ((uint64_t(1) << b->sort_bits) < P + 1 && b->sort_bits < 32) { b->sort_bits++; }

    cub::DeviceRadixSort::SortPairs(nullptr, b->sort_temp_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                    static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<int>(Z), 0, b->sort_bits, ctx->stream);

    b->rowsort_bits = 33;
    while ((uint64_t(1) << (b->rowsort_bits - 32)) < C && b->rowsort_bits < 64) { b->rowsort_bits++; }

    cub::DeviceRadixSort::SortPairs(nullptr, b->rowsort_temp_bytes, static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr),
                                    static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<int>(R), 0, b->rowsort_bits, ctx->stream);
    cub::DeviceScan::ExclusiveSum(nullptr, b->scan_temp_bytes, static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr), static_cast<int>(R + 1), ctx->stream);

    size_t bytes = 0;
    bytes += 4 * padded<uint64_t>(R + 1) + 2 * padded<uint32_t>(R) + padded<uint32_t>(Z) + padded<double>(R);
    bytes += ((b->rowsort_temp_bytes + 255) & ~size_t(255)) + ((b->scan_temp_bytes + 255) & ~size_t(255));
    bytes += 2 * padded<uint64_t>(C + 1) + padded<uint64_t>(R + 1) + padded<uint32_t>(Z) + padded<double>(Z) + padded<double>(R) + padded<uint32_t>(R) + padded<uint32_t>(C);
    bytes += 2 * padded<uint64_t>(C);
    bytes += padded<double>(Z) + padded<double>(R) + 5 * padded<uint32_t>(Z) + padded<uint64_t>(P + 1) + padded<uint32_t>(Z) + padded<double>(Z);
    bytes += ((b->sort_temp_bytes + 255) & ~size_t(255));
    bytes += padded<double>(P + C) + padded<double>(R) + padded<double>(P) + 2 * padded<double>(C) + padded<uint32_t>(C) + padded<uint32_t>(1);
    bytes += ((b->scratch_bytes + 255) & ~size_t(255));
    bytes += padded<double>(R) + 8192;
    bytes += padded<double>(2 * kStreamMaxCtas) + padded<uint32_t>(kStreamMaxCtas);

    cudaError_t err = cudaMallocFromPoolAsync(reinterpret_cast<void **>(&b->pool.base), bytes, ctx->pool, ctx->stream);
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
    b->d_count_f64 = pool.take<double>(R);
    b->d_order = pool.take<uint32_t>(C);
    b->d_group_scratch = pool.take<uint64_t>(C);
    b->d_group_cta_bytes = pool.take<uint64_t>(C);
    b->d_rowkey_in = pool.take<uint64_t>(R + 1);
    b->d_rowkey_out = pool.take<uint64_t>(R + 1);
    b->d_rowid_in = pool.take<uint32_t>(R);
    b->d_perm = pool.take<uint32_t>(R);
    b->d_sorted_len = pool.take<uint64_t>(R + 1);
    b->d_sorted_row_ptr = pool.take<uint64_t>(R + 1);
    b->d_sorted_col = pool.take<uint32_t>(Z);
    b->d_sorted_noise = pool.take<double>(R);
    b->d_rowsort_temp = pool.take<char>(b->rowsort_temp_bytes);
    b->d_scan_temp = pool.take<char>(b->scan_temp_bytes);
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
    b->d_scratch = pool.take<unsigned char>(b->scratch_bytes);
    b->d_stream_partial = pool.take<double>(2 * kStreamMaxCtas);
    b->d_stream_flags = pool.take<uint32_t>(kStreamMaxCtas);

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

    RPVG_UPLOAD(b->d_read_offset, read_offset, C + 1, uint64_t)
    RPVG_UPLOAD(b->d_path_offset, path_offset, C + 1, uint64_t)
    // the plan vectors are temporaries: the synchronize below keeps them alive long enough
    RPVG_UPLOAD(b->d_order, order.data(), C, uint32_t)
    RPVG_UPLOAD(b->d_group_scratch, group_scratch.data(), C, uint64_t)
    RPVG_UPLOAD(b->d_group_cta_bytes, group_cta_bytes.data(), C, uint64_t)
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

    int status = rpvgCheckClusters(ctx, cl);
    if (status != RPVG_B200_OK) { return status; }

    std::vector<uint64_t> cluster_nnz(cl->n_clusters);

    for (uint32_t c = 0; c < cl->n_clusters; ++c) {
        const uint64_t lo = cl->row_ptr[cl->read_offset[c]];
        const uint64_t hi = cl->row_ptr[cl->read_offset[c + 1]];
        if (hi < lo) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "row_ptr must be non-decreasing"); }
        cluster_nnz[c] = hi - lo;
    }

    rpvg_b200_batch * b = nullptr;
    status = rpvgBatchCreate(ctx, cl->n_clusters, cl->read_offset, cl->path_offset, cluster_nnz.data(), false, &b);
    if (status != RPVG_B200_OK) { return status; }

    const uint64_t R = b->n_reads, Z = b->nnz;
    cudaStream_t s = ctx->stream;
    cudaError_t err = cudaSuccess;

    if (err == cudaSuccess) { err = cudaMemcpyAsync(b->d_row_ptr, cl->row_ptr, (R + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, s); }
    if (err == cudaSuccess && Z > 0) { err = cudaMemcpyAsync(b->d_col, cl->col, Z * sizeof(uint32_t), cudaMemcpyHostToDevice, s); }
    if (err == cudaSuccess && Z > 0) { err = cudaMemcpyAsync(b->d_prob, cl->prob, Z * sizeof(double), cudaMemcpyHostToDevice, s); }
    if (err == cudaSuccess && R > 0) { err = cudaMemcpyAsync(b->d_noise, cl->noise, R * sizeof(double), cudaMemcpyHostToDevice, s); }
    if (err == cudaSuccess && R > 0) { err = cudaMemcpyAsync(b->d_count, cl->count, R * sizeof(uint32_t), cudaMemcpyHostToDevice, s); }
    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

    if (err != cudaSuccess) {
        rpvg_b200_batch_free(ctx, b);
        return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("arena upload: ") + cudaGetErrorString(err));
    }

    *out_batch = b;
    return RPVG_B200_OK;
}

static int batchEmLaunch(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, const rpvg_b200_em_params * params);

int rpvg_b200_batch_em(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, const rpvg_b200_em_params * params) {

    const int status = batchEmLaunch(ctx, b, params);

    // A failure between the fork and the join leaves side and lane streams with work the ctx stream does not wait for: the
    // caller may free the batch next, so everything is drained here before the error goes back.
    if (status != RPVG_B200_OK && ctx) {
        const std::string message = ctx->error;
        for (auto side: ctx->side_streams) { cudaStreamSynchronize(side); }
        for (auto lane: ctx->lane_streams) { cudaStreamSynchronize(lane); }
        cudaStreamSynchronize(ctx->stream);
        cudaGetLastError();
        ctx->error = message;
    }

    return status;
}

static int batchEmLaunch(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, const rpvg_b200_em_params * params) {

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
        const unsigned grid1 = static_cast<unsigned>((b->n_reads + 1 + 255) / 256);

        rowKeysKernel<<<grid, 256, 0, s>>>(b->n_clusters, b->n_reads, b->d_read_offset, b->d_row_ptr, b->d_col, b->d_rowkey_in, b->d_rowid_in);
        RPVG_LAUNCHED(ctx);

        size_t temp_bytes = b->rowsort_temp_bytes;
        RPVG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(b->d_rowsort_temp, temp_bytes, b->d_rowkey_in, b->d_rowkey_out, b->d_rowid_in, b->d_perm,
                                                       static_cast<int>(b->n_reads), 0, b->rowsort_bits, s));
        ctx->launches += 2 + static_cast<uint64_t>((b->rowsort_bits + 7) / 8);

        sortedLengthsKernel<<<grid1, 256, 0, s>>>(b->n_reads, b->d_perm, b->d_row_ptr, b->d_sorted_len);
        RPVG_LAUNCHED(ctx);

        temp_bytes = b->scan_temp_bytes;
        RPVG_CUDA(ctx, cub::DeviceScan::ExclusiveSum(b->d_scan_temp, temp_bytes, b->d_sorted_len, b->d_sorted_row_ptr, static_cast<int>(b->n_reads + 1), s));
        ctx->launches += 2;

        buildRowsKernel<<<grid, 256, 0, s>>>(b->n_clusters, b->n_reads, b->d_read_offset, b->d_path_offset, b->d_row_ptr, b->d_col, b->d_prob,
                                             b->d_noise, b->d_count, b->d_perm, b->d_sorted_row_ptr, b->d_sorted_col, b->d_sorted_noise,
                                             b->d_val, b->d_cnt, b->d_entry_row, b->d_key_in, b->d_idx_in, b->d_error, b->d_count_f64, b->prenormalised ? 1 : 0);
        RPVG_LAUNCHED(ctx);

    } else {

        RPVG_CUDA(ctx, cudaMemsetAsync(b->d_sorted_row_ptr, 0, sizeof(uint64_t), s));
    }

    if (b->nnz > 0) {

        size_t temp_bytes = b->sort_temp_bytes;
        RPVG_CUDA(ctx, cub::DeviceRadixSort::SortPairs(b->d_sort_temp, temp_bytes, b->d_key_in, b->d_key_out, b->d_idx_in, b->d_idx_out,
                                                       static_cast<int>(b->nnz), 0, b->sort_bits, s));
        ctx->launches += 2 + static_cast<uint64_t>((b->sort_bits + 7) / 8);   // histogram, scan and one pass per 8 key bits
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
    args.row_ptr = b->d_sorted_row_ptr;
    args.col = b->d_sorted_col;
    args.val = b->d_val;
    args.noise = b->d_sorted_noise;
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
    args.error = b->d_error;
    args.params = *params;

    RPVG_CUDA(ctx, cudaEventRecord(b->ev_built, s));

    // Giant clusters first, one after the other on the ctx stream: each is a cooperative launch that fills the GPU.
    for (size_t k = 0; k < b->bins.size(); ++k) {

        const auto & bin = b->bins[k];
        if (!bin.stream) { continue; }

        for (uint32_t i = 0; i < bin.count; ++i) {

            StreamArgs sargs;
            sargs.em = args;
            sargs.cluster = b->h_order[bin.first + i];
            sargs.partial = b->d_stream_partial;
            sargs.low_partial = b->d_stream_partial + kStreamMaxCtas;
            sargs.flags = b->d_stream_flags;

            void * kernel_args[] = {&sargs};
            RPVG_CUDA(ctx, cudaLaunchCooperativeKernel(reinterpret_cast<void *>(emStreamKernel), dim3(ctx->stream_grid), dim3(kStreamThreads), kernel_args, 0, s));
            RPVG_LAUNCHED(ctx);
        }
    }

    // Bins run concurrently on the ctx's side streams (fork after the build, join before ev_done), one stream priority
    // level per bin, larger groups higher: a thread-block cluster of 16 needs 16 idle SMs of one GPC at the same moment,
    // which it only gets if it is placed before anything smaller. Measured on B200 (tools/em_timeline.py): 7 of the 8
    // GPCs can host a cluster of 16; what they leave over (2-4 SMs each, 10 in the small GPC) is filled by the bins
    // of groups of 2 and 8 while the big ones run.
    RPVG_CUDA(ctx, cudaEventRecord(ctx->fork_event, s));

    const bool debug = std::getenv("RPVG_B200_DEBUG") != nullptr;
    const uint32_t spacer_us = envKnob("RPVG_B200_EM_SPACER_US", 20);
    const uint32_t lane_min_group = envKnob("RPVG_B200_EM_LANE_MIN_GROUP", 16);
    const size_t n_lanes = std::min<size_t>(envKnob("RPVG_B200_EM_LANES", 7), ctx->lane_streams.size());
    // Bin -> side stream. Priority classes, highest first (the groups of 16 run on the lane streams, above all):
    // groups of 2, groups of 4, single CTAs with a full shared-memory tier, groups of 8, everything small. Small
    // footprints first: they fit the SMs the big clusters leave over in each GPC (measured with tools/em_timeline.py: a
    // group of 8 pending at a higher level keeps those SMs idle instead); the small single-CTA launches are the filler
    // at the end. RPVG_B200_EM_STREAMS="i j k ..." overrides the choice for tuning runs.
    std::vector<size_t> stream_map;
    if (const char * v = std::getenv("RPVG_B200_EM_STREAMS")) {
        for (const char * q = v; *q;) {
            char * end = nullptr;
            const unsigned long x = std::strtoul(q, &end, 10);
            if (end == q) { break; }
            stream_map.push_back(x % ctx->side_streams.size());
            q = end;
        }
    }

    size_t class_cursor[5] = {0, 0, 0, 0, 0};
    auto stream_of_bin = [&](size_t k) -> size_t {
        if (k < stream_map.size()) { return stream_map[k]; }
        const auto & bin = b->bins[k];
        const uint32_t top_smem = std::min<uint32_t>(sellTiers().smem[kNumTiers - 1], static_cast<uint32_t>(ctx->max_smem_optin) - kStaticSmemReserve);
        // the dense kernels share the last stream of the lowest class: their CTAs are many and small, a launch of
        // theirs queued behind a sliced-ELL bin would start when that bin has drained
        if (bin.dense_dim > 0) { return ctx->side_streams.size() - 1; }
        size_t cls = 4;
        if (bin.group_size == 2) { cls = 0; }
        else if (bin.group_size == 4) { cls = 1; }
        else if (bin.group_size == 1 && bin.smem >= top_smem) { cls = 2; }
        else if (bin.group_size >= 8) { cls = 3; }
        const size_t width = cls == 4 ? 3 : 2;
        return 2 * cls + (class_cursor[cls]++ % width);
    };

    std::vector<double> lane_load(n_lanes, 0.0);
    std::vector<bool> lane_forked(n_lanes, false);
    bool lanes_used = false;

    for (size_t k = 0; k < b->bins.size(); ++k) {

        const auto & bin = b->bins[k];
        if (debug) {
            fprintf(stderr, "[rpvg_b200] em bin %zu: group %u threads %u smem %u index %u clusters %u (max group %u, scratch %.1f MB)\n", k, bin.group_size,
                    bin.threads, bin.smem, bin.index_bytes, bin.count, ctx->max_group_size, b->scratch_bytes / 1e6);
        }

        if (bin.stream) { continue; }   // done above

        cudaStream_t bs = ctx->side_streams[stream_of_bin(k)];
        RPVG_CUDA(ctx, cudaStreamWaitEvent(bs, ctx->fork_event, 0));

        // The launch of the largest groups goes first and alone: its thread-block clusters need whole GPCs, which they
        // only find while the GPU is empty. Everything else starts a few microseconds later.
        if (spacer_us > 0 && k > 0 && b->bins[0].group_size >= 8) {
            spacerKernel<<<1, 1, 0, bs>>>(1000ull * spacer_us);
            RPVG_LAUNCHED(ctx);
        }

        if (bin.dense_dim > 0) {

            const cudaError_t launch_err = launchDenseBin(args, b->d_order + bin.first, bin, bs);
            if (launch_err != cudaSuccess) {
                return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("emDenseKernel launch (side ") + std::to_string(bin.dense_dim) + "): " + cudaGetErrorString(launch_err));
            }
            RPVG_LAUNCHED(ctx);

        } else if (bin.group_size == 0) {

            EmArgs fallback_args = args;
            fallback_args.order = b->d_order + bin.first;
            emClusterKernel<<<bin.count, kEmThreads, 0, bs>>>(fallback_args);
            RPVG_LAUNCHED(ctx);

        } else {

            SellArgs sa;
            sa.em = args;
            sa.group_cluster = b->d_order + bin.first;
            sa.group_scratch = b->d_group_scratch + bin.first;
            sa.group_cta_bytes = b->d_group_cta_bytes + bin.first;
            sa.scratch = b->d_scratch;
            sa.smem_bytes = bin.smem;

            cudaError_t launch_err = cudaSuccess;

            if (bin.group_size >= lane_min_group && n_lanes > 0) {

                // One launch per cluster, dealt over the lane streams: a launch with more thread-block clusters than fit
                // at once would keep the SMs it cannot use idle until its last cluster is placed (the work distributor
                // holds freed SMs back for the pending clusters of the kernel it is dispatching).
                // (within a bin the clusters come heaviest first: each goes to the lane with the least expected work,
                // cost ~ weight^1.5, i.e. entries x iterations)
                for (uint32_t i = 0; i < bin.count && launch_err == cudaSuccess; ++i) {

                    size_t lane = 0;
                    for (size_t l = 1; l < n_lanes; ++l) {
                        if (lane_load[l] < lane_load[lane]) { lane = l; }
                    }

                    const uint32_t c = b->h_order[bin.first + i];
                    const double weight = static_cast<double>(b->h_cluster_nnz[c] + (b->h_read_offset[c + 1] - b->h_read_offset[c]));
                    lane_load[lane] += weight * std::sqrt(weight);

                    if (!lane_forked[lane]) {
                        RPVG_CUDA(ctx, cudaStreamWaitEvent(ctx->lane_streams[lane], ctx->fork_event, 0));
                        lane_forked[lane] = true;
                    }

                    SellArgs one = sa;
                    one.group_cluster += i;
                    one.group_scratch += i;
                    one.group_cta_bytes += i;

                    rpvg_b200_batch::EmBin single = bin;
                    single.count = 1;
                    launch_err = launchSellBin(one, single, ctx->lane_streams[lane]);
                    if (i > 0) { RPVG_LAUNCHED(ctx); RPVG_LAUNCHED(ctx); }
                }

                lanes_used = true;

            } else {
                launch_err = launchSellBin(sa, bin, bs);
            }

            if (launch_err != cudaSuccess) {
                return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("emSell kernels launch (group ") + std::to_string(bin.group_size) + ", threads " +
                                std::to_string(bin.threads) + ", smem " + std::to_string(bin.smem) + "): " + cudaGetErrorString(launch_err));
            }
            RPVG_LAUNCHED(ctx);
            RPVG_LAUNCHED(ctx);   // build + run
        }

    }

    for (size_t k = 0; k < ctx->side_streams.size(); ++k) {   // (streams no bin went to join at once)
        RPVG_CUDA(ctx, cudaEventRecord(ctx->join_events[k], ctx->side_streams[k]));
        RPVG_CUDA(ctx, cudaStreamWaitEvent(s, ctx->join_events[k], 0));
    }

    for (size_t k = 0; lanes_used && k < n_lanes; ++k) {
        if (!lane_forked[k]) { continue; }
        RPVG_CUDA(ctx, cudaEventRecord(ctx->lane_events[k], ctx->lane_streams[k]));
        RPVG_CUDA(ctx, cudaStreamWaitEvent(s, ctx->lane_events[k], 0));
    }

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

    if (dev_error & RPVG_DEV_ERR_CAPACITY) {
        return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "internal: a cluster slice outgrew its scratch area (please report; RPVG_B200_EM_SLICE_WEIGHT changes the slicing)");
    }

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

    return rpvg_b200_em_abundances_gibbs(ctx, clusters, params, nullptr, abundances, noise_count, total_count, iterations);
}

int rpvg_b200_em_abundances_gibbs(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_em_params * params, const rpvg_b200_gibbs * gibbs,
                                  double * abundances, double * noise_count, double * total_count, uint32_t * iterations) {

    rpvg_b200_batch * batch = nullptr;

    int status = rpvg_b200_batch_upload(ctx, clusters, &batch);
    if (status != RPVG_B200_OK) { return status; }

    status = rpvg_b200_batch_em(ctx, batch, params);
    if (status == RPVG_B200_OK) { status = rpvg_b200_batch_download(ctx, batch, abundances, noise_count, total_count, iterations); }

    if (status == RPVG_B200_OK && gibbs) {
        status = rpvg_b200_batch_gibbs(ctx, batch, gibbs->num_samples, gibbs->thin_its, gibbs->seeds, gibbs->abundance_samples, gibbs->noise_samples);
    }

    rpvg_b200_batch_free(ctx, batch);
    return status;
}

}

#ifdef RPVG_B200_EM_TIMELINE
// tuning builds only: copies the per-CTA records out (12 values each) and resets the counter
extern "C" int rpvg_b200_timeline_read(unsigned long long * records, unsigned int capacity, unsigned int * count) {

    unsigned int n = 0;
    if (cudaDeviceSynchronize() != cudaSuccess) { return -1; }
    if (cudaMemcpyFromSymbol(&n, rpvg_timeline_count, sizeof(n)) != cudaSuccess) { return -1; }
    n = n < 65536 ? n : 65536;
    n = n < capacity ? n : capacity;
    if (n > 0 && cudaMemcpyFromSymbol(records, rpvg_timeline_records, 12ull * 8 * n) != cudaSuccess) { return -1; }
    const unsigned int zero = 0;
    if (cudaMemcpyToSymbol(rpvg_timeline_count, &zero, sizeof(zero)) != cudaSuccess) { return -1; }
    *count = n;
    return 0;
}
#endif
