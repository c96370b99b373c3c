// rpvg_b200_add_path_probs — the per-read path probabilities, the step right before the estimators' boundary.
//
// Reference semantics restated here (never its code): ReadPathProbabilities::addPathProbs,
// rpvg:src/read_path_probabilities.cpp:74-221, with calcAlignPathLogProbs (:39-67):
//   noise = max(precision, min_noise, 10^(-mapq/10)); noise += (1 - noise) exp(noise alignment log-probability);
//   per path the log-probability of the alignment path with the longest alignment (ties: the largest value), minus
//   log(effective length); optionally summed per path group (add_log of value + log(source_count));
//   normalised with a running add_log over ascending index; probabilities below the precision go to the noise;
//   the others are grouped greedily (first group whose running mean lies within the precision), scaled by (1 - noise)
//   and the groups sorted by (probability, ids).
//
// The reference keeps two dense vectors over all paths of the cluster per read (O(N) per read). Here a read touches only
// the paths its alignment paths list: (read, path) tuples are sorted once for the whole batch (radix sort), duplicates
// reduced, and one thread per read walks its own few entries. Paths a read does not touch hold the most negative double
// in the reference and add exactly nothing to its add_log chains (exp underflows to 0), so the sparse walk over
// ascending indices is the same chain.
//
// exp/log/log1p are the device's (<= 1-2 ulp from glibc's): results agree with the reference to ~1e-14 relative, the
// grouping and the order are identical unless a value sits within that distance of a threshold.

#include <algorithm>
#include <cfloat>
#include <cmath>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "em_batch.cuh"

namespace {

constexpr double kScoreLogBase = 1.383325268738;   // rpvg:src/utils.hpp:83
constexpr double kNoiseScoreLogBase = 1e-6;        // rpvg:src/utils.hpp:84
constexpr double kLowest = -DBL_MAX;

enum : uint32_t {
    PATHPROBS_ERR_PATH_RANGE = 1u,   // a path index >= the number of paths of the read's cluster
    PATHPROBS_ERR_NO_PATH = 2u,      // a read with a non-zero noise score and no path of positive effective length
    PATHPROBS_ERR_LAYOUT = 4u        // a read without alignment paths, or ids on its noise alignment path
};

struct Alignments {

    uint32_t n_clusters;
    uint64_t n_reads, n_align, n_ids;
    const uint64_t * read_offset;
    const uint64_t * path_offset;
    const double * effective_length;
    const uint32_t * path_group;     // or NULL
    const uint32_t * path_rank;      // position of each path in (group, path) order within its cluster; NULL = the path index
    const uint32_t * source_count;   // or NULL
    const uint8_t * min_mapq;
    const uint64_t * align_ptr;
    const int32_t * score_sum;
    const uint32_t * align_length;
    const double * frag_log_prob;    // or NULL
    const uint64_t * id_ptr;
    const uint32_t * ids;
};

// first i with ptr[i + 1] > x (ptr non-decreasing, ptr[n] > x)
__device__ __forceinline__ uint64_t segmentOf(const uint64_t * __restrict__ ptr, uint64_t n, uint64_t x) {

    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid + 1) > x) { hi = mid; } else { lo = mid + 1; }
    }
    return lo;
}

// rpvg:src/utils.hpp:300-302
__device__ __forceinline__ double addLog(double log_x, double log_y) {

    return log_x > log_y ? log_x + log1p(exp(log_y - log_x)) : log_y + log1p(exp(log_x - log_y));
}

// One thread per (alignment path, path id): key = (read, rank of the path), payload = the tuple's index.
// rpvg:src/read_path_probabilities.cpp:39-67 (log-probability of the alignment path) and :128 (minus log effective length).
__global__ void pathTupleKernel(const Alignments al, uint64_t * __restrict__ key, uint32_t * __restrict__ tuple, double * __restrict__ tuple_lp,
                                uint32_t * __restrict__ tuple_len, uint32_t * __restrict__ error) {

    const uint64_t t = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (t >= al.n_ids) { return; }

    const uint64_t a = segmentOf(al.id_ptr, al.n_align, t);
    const uint64_t r = segmentOf(al.align_ptr, al.n_reads, a);
    const uint32_t c = clusterOf(al.read_offset, al.n_clusters, r);

    const uint64_t p0 = al.path_offset[c];
    const uint64_t N = al.path_offset[c + 1] - p0;
    const uint32_t path = al.ids[t];

    tuple[t] = static_cast<uint32_t>(t);

    if (a + 1 == al.align_ptr[r + 1]) {   // the noise alignment path lists no paths
        atomicOr(error, PATHPROBS_ERR_LAYOUT);
        key[t] = (r << 32) | 0xffffffffu;
        return;
    }

    if (path >= N) {
        atomicOr(error, PATHPROBS_ERR_PATH_RANGE);
        key[t] = (r << 32) | 0xffffffffu;
        return;
    }

    const double eff = al.effective_length[p0 + path];

    if (eff == 0) {   // Utils::doubleCompare(effective_length, 0): the path keeps the most negative value
        key[t] = (r << 32) | 0xffffffffu;
        return;
    }

    double lp = static_cast<double>(al.score_sum[a]) * kScoreLogBase;
    if (al.frag_log_prob) { lp += al.frag_log_prob[a]; }

    tuple_lp[t] = lp - log(eff);
    tuple_len[t] = al.align_length[a];

    const uint32_t rank = al.path_rank ? al.path_rank[p0 + path] : path;
    key[t] = (r << 32) | rank;
}

// One thread per sorted tuple: the first of a run of equal (read, path) keeps the run's winner — longest alignment, then
// largest value (rpvg:src/read_path_probabilities.cpp:131-142 applied in any order gives the lexicographic maximum).
__global__ void pathHeadKernel(uint64_t n, const uint64_t * __restrict__ key, const uint32_t * __restrict__ tuple, const double * __restrict__ tuple_lp,
                               const uint32_t * __restrict__ tuple_len, uint32_t * __restrict__ head, double * __restrict__ head_lp) {

    const uint64_t k = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (k > n) { return; }

    if (k == n) {   // one extra slot so that the scan yields the total
        head[k] = 0;
        return;
    }

    const uint64_t kk = key[k];
    const bool valid = static_cast<uint32_t>(kk) != 0xffffffffu;
    const bool first = valid && (k == 0 || key[k - 1] != kk);

    head[k] = first ? 1u : 0u;
    if (!first) { return; }

    uint32_t best_len = tuple_len[tuple[k]];
    double best_lp = tuple_lp[tuple[k]];

    for (uint64_t q = k + 1; q < n && key[q] == kk; ++q) {
        const uint32_t len = tuple_len[tuple[q]];
        const double lp = tuple_lp[tuple[q]];
        if (len > best_len || (len == best_len && lp > best_lp)) {
            best_len = len;
            best_lp = lp;
        }
    }

    head_lp[k] = best_lp;
}

__global__ void pathCompactKernel(uint64_t n, const uint64_t * __restrict__ key, const uint32_t * __restrict__ head, const uint32_t * __restrict__ head_pos,
                                  const double * __restrict__ head_lp, uint32_t * __restrict__ entry_rank, double * __restrict__ entry_lp) {

    const uint64_t k = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (k >= n || !head[k]) { return; }

    const uint32_t u = head_pos[k];
    entry_rank[u] = static_cast<uint32_t>(key[k]);
    entry_lp[u] = head_lp[k];
}

// One thread per read: everything that is sequential in the reference (rpvg:src/read_path_probabilities.cpp:94-219) over
// the read's own entries [u0, u1), ascending by index. Work arrays are indexed like the entries.
__global__ void pathReadKernel(const Alignments al, const double min_noise_prob, const double prob_precision, const double * __restrict__ phred_prob,
                               const uint32_t * __restrict__ head_pos, const uint32_t * __restrict__ rank_path, uint32_t * __restrict__ entry_rank,
                               double * __restrict__ entry_lp, uint32_t * __restrict__ member, double * __restrict__ group_first, uint32_t * __restrict__ group_size,
                               uint32_t * __restrict__ group_start, uint32_t * __restrict__ group_order, uint32_t * __restrict__ staged_ids, double * __restrict__ noise,
                               uint32_t * __restrict__ n_groups, uint32_t * __restrict__ n_out_ids, uint32_t * __restrict__ error) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= al.n_reads) { return; }

    n_groups[r] = 0;
    n_out_ids[r] = 0;

    const uint64_t a0 = al.align_ptr[r], a1 = al.align_ptr[r + 1];

    if (a1 < a0 + 2) {   // rpvg asserts align_paths.size() > 1
        atomicOr(error, PATHPROBS_ERR_LAYOUT);
        noise[r] = 1;
        return;
    }

    const uint32_t mapq = al.min_mapq[r];

    if (mapq == 0) {   // rpvg:src/read_path_probabilities.cpp:87 — nothing is added, the record keeps noise 1
        noise[r] = 1;
        return;
    }

    // :92, :100
    double noise_prob = fmax(prob_precision, fmax(min_noise_prob, phred_prob[mapq]));
    noise_prob += (1 - noise_prob) * exp(static_cast<double>(al.score_sum[a1 - 1]) * kNoiseScoreLogBase);

    if (al.score_sum[a1 - 1] == 0) {   // :102-106
        noise[r] = noise_prob;
        return;
    }

    const uint32_t c = clusterOf(al.read_offset, al.n_clusters, r);
    const uint64_t p0 = al.path_offset[c];

    const uint32_t u0 = head_pos[al.id_ptr[a0]];
    const uint32_t u1 = head_pos[al.id_ptr[a1]];

    // entries -> (index, value): the path itself, or its group with the add_log chain over the group's paths (:147-163)
    uint32_t n = 0;

    if (al.path_group) {

        for (uint32_t k = u0; k < u1;) {

            const uint32_t path = rank_path[p0 + entry_rank[k]];
            const uint32_t group = al.path_group[p0 + path];
            double v = kLowest;

            while (k < u1) {
                const uint32_t q = rank_path[p0 + entry_rank[k]];
                if (al.path_group[p0 + q] != group) { break; }
                v = addLog(v, entry_lp[k] + log(static_cast<double>(al.source_count ? al.source_count[p0 + q] : 1u)));
                ++k;
            }

            entry_rank[u0 + n] = group;
            entry_lp[u0 + n] = v;
            ++n;
        }

    } else {

        n = u1 - u0;
    }

    // :165-170
    double sum = kLowest;
    for (uint32_t k = 0; k < n; ++k) { sum = addLog(sum, entry_lp[u0 + k]); }

    if (!(sum > kLowest)) {   // rpvg asserts
        atomicOr(error, PATHPROBS_ERR_NO_PATH);
        noise[r] = noise_prob;
        return;
    }

    // :176-210
    double low_prob_sum = 0;
    uint32_t G = 0;

    for (uint32_t k = 0; k < n; ++k) {

        const double p = exp(entry_lp[u0 + k] - sum);
        uint32_t g = 0xffffffffu;

        if (p >= prob_precision) {

            for (g = 0; g < G; ++g) {
                if (fabs(group_first[u0 + g] - p) < prob_precision) {
                    const double size = static_cast<double>(group_size[u0 + g]);
                    group_first[u0 + g] = (group_first[u0 + g] * size + p) / (size + 1);
                    group_size[u0 + g]++;
                    break;
                }
            }

            if (g == G) {
                group_first[u0 + G] = p;
                group_size[u0 + G] = 1;
                ++G;
            }

        } else {

            low_prob_sum += p;
        }

        member[u0 + k] = g;
    }

    // :212-217
    for (uint32_t g = 0; g < G; ++g) { group_first[u0 + g] *= (1 - noise_prob); }
    noise_prob += low_prob_sum * (1 - noise_prob);
    noise[r] = noise_prob;

    // the ids of every group, ascending (entries are ascending), group after group in creation order
    uint32_t total = 0;
    for (uint32_t g = 0; g < G; ++g) {
        group_start[u0 + g] = total;
        total += group_size[u0 + g];
        group_size[u0 + g] = 0;   // refilled as a cursor
    }
    for (uint32_t k = 0; k < n; ++k) {
        const uint32_t g = member[u0 + k];
        if (g != 0xffffffffu) { staged_ids[u0 + group_start[u0 + g] + group_size[u0 + g]++] = entry_rank[u0 + k]; }
    }

    // :219 — sort by (probability, ids): insertion sort over the few groups of a read
    auto less = [&](uint32_t x, uint32_t y) {

        const double fx = group_first[u0 + x], fy = group_first[u0 + y];
        if (fx != fy) { return fx < fy; }

        const uint32_t sx = group_size[u0 + x], sy = group_size[u0 + y];
        const uint32_t * ix = staged_ids + u0 + group_start[u0 + x];
        const uint32_t * iy = staged_ids + u0 + group_start[u0 + y];

        for (uint32_t q = 0; q < min(sx, sy); ++q) {
            if (ix[q] != iy[q]) { return ix[q] < iy[q]; }
        }
        return sx < sy;
    };

    for (uint32_t g = 0; g < G; ++g) {
        uint32_t q = g;
        while (q > 0 && less(g, group_order[u0 + q - 1])) {
            group_order[u0 + q] = group_order[u0 + q - 1];
            --q;
        }
        group_order[u0 + q] = g;
    }

    n_groups[r] = G;
    n_out_ids[r] = total;
}

// One thread per read: its groups, in sorted order, to their final places.
__global__ void pathEmitKernel(const Alignments al, const uint32_t * __restrict__ head_pos, const double * __restrict__ group_first, const uint32_t * __restrict__ group_size,
                               const uint32_t * __restrict__ group_start, const uint32_t * __restrict__ group_order, const uint32_t * __restrict__ staged_ids,
                               const uint32_t * __restrict__ n_groups, const uint64_t * __restrict__ group_base, const uint64_t * __restrict__ id_base,
                               uint64_t * __restrict__ out_group_ptr, double * __restrict__ out_group_prob, uint64_t * __restrict__ out_id_ptr, uint32_t * __restrict__ out_ids) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r > al.n_reads) { return; }

    out_group_ptr[r] = group_base[r];

    if (r == al.n_reads) {
        out_id_ptr[group_base[r]] = id_base[r];
        return;
    }

    const uint32_t G = n_groups[r];
    if (G == 0) { return; }

    const uint32_t u0 = head_pos[al.id_ptr[al.align_ptr[r]]];
    uint64_t out_group = group_base[r];
    uint64_t out_id = id_base[r];

    for (uint32_t k = 0; k < G; ++k) {

        const uint32_t g = group_order[u0 + k];
        const uint32_t size = group_size[u0 + g];

        out_group_prob[out_group] = group_first[u0 + g];
        out_id_ptr[out_group] = out_id;
        ++out_group;

        for (uint32_t q = 0; q < size; ++q) { out_ids[out_id++] = staged_ids[u0 + group_start[u0 + g] + q]; }
    }
}

}

extern "C" int rpvg_b200_add_path_probs(rpvg_b200_ctx * ctx, const rpvg_b200_alignments * in, double min_noise_prob, double prob_precision, double * noise,
                                        uint64_t * group_ptr, double * group_prob, uint64_t * group_id_ptr, uint32_t * group_ids) {

    if (!ctx || !in) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: NULL argument"); }
    if (in->n_clusters == 0) { return RPVG_B200_OK; }
    if (!in->read_offset || !in->path_offset || !in->effective_length || !group_ptr) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: NULL array"); }

    const uint32_t C = in->n_clusters;
    const uint64_t R = in->read_offset[C];
    const uint64_t P = in->path_offset[C];

    if (R == 0) {
        group_ptr[0] = 0;
        if (group_id_ptr) { group_id_ptr[0] = 0; }
        return RPVG_B200_OK;
    }

    if (!in->min_mapq || !in->align_ptr || !in->score_sum || !in->align_length || !in->id_ptr || !noise || !group_prob || !group_id_ptr || !group_ids) {
        return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: NULL array");
    }

    for (uint32_t c = 0; c < C; ++c) {
        if (in->read_offset[c + 1] < in->read_offset[c] || in->path_offset[c + 1] < in->path_offset[c]) {
            return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: offsets must be non-decreasing");
        }
        if (in->path_offset[c + 1] - in->path_offset[c] >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_add_path_probs: too many paths in a cluster"); }
    }

    const uint64_t A = in->align_ptr[R];
    const uint64_t T = in->id_ptr[A];

    if (T > 0 && !in->ids) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: ids is NULL"); }
    if (R >= (1ull << 31) || T >= (1ull << 31) || A >= 0xffffffffull) {
        return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_add_path_probs: reads and path ids must each stay below 2^31 per call");
    }

    for (uint64_t r = 0; r < R; ++r) {
        if (in->align_ptr[r + 1] < in->align_ptr[r]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: align_ptr must be non-decreasing"); }
    }
    for (uint64_t a = 0; a < A; ++a) {
        if (in->id_ptr[a + 1] < in->id_ptr[a]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_add_path_probs: id_ptr must be non-decreasing"); }
    }

    // Utils::phred_to_prob (rpvg:src/utils.hpp:131-133) for every possible mapping quality
    double phred_prob[256];
    for (int q = 0; q < 256; ++q) { phred_prob[q] = std::pow(10, -static_cast<double>(q) / 10); }

    // collapse_groups: the chains run group by group, the paths of a group ascending — rank of each path in that order
    std::vector<uint32_t> path_rank, rank_path;
    const bool collapse = in->path_group != nullptr;

    if (collapse) {

        path_rank.resize(P);
        rank_path.resize(P);

        for (uint32_t c = 0; c < C; ++c) {

            const uint64_t p0 = in->path_offset[c], p1 = in->path_offset[c + 1];
            for (uint64_t p = p0; p < p1; ++p) { rank_path[p] = static_cast<uint32_t>(p - p0); }

            std::stable_sort(rank_path.begin() + p0, rank_path.begin() + p1, [&](uint32_t x, uint32_t y) { return in->path_group[p0 + x] < in->path_group[p0 + y]; });
            for (uint64_t k = p0; k < p1; ++k) { path_rank[p0 + rank_path[k]] = static_cast<uint32_t>(k - p0); }
        }
    }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    int read_bits = 1;
    while ((1ull << read_bits) < R && read_bits < 32) { read_bits++; }

    size_t sort_bytes = 0, scan_bytes = 0, scan64_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr), static_cast<uint32_t *>(nullptr),
                                    static_cast<uint32_t *>(nullptr), static_cast<int>(T), 0, 32 + read_bits, s);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint32_t *>(nullptr), static_cast<int>(T + 1), s);
    cub::DeviceScan::ExclusiveSum(nullptr, scan64_bytes, static_cast<uint32_t *>(nullptr), static_cast<uint64_t *>(nullptr), static_cast<int>(R + 1), s);

    const size_t bytes = 2 * padded<uint64_t>(C + 1) + padded<double>(P) + 4 * padded<uint32_t>(P) + padded<uint8_t>(R) + padded<uint64_t>(R + 1) + padded<int32_t>(A) +
                         padded<uint32_t>(A) + padded<double>(A) + padded<uint64_t>(A + 1) + padded<uint32_t>(T + 1) + padded<double>(256) +
                         2 * padded<uint64_t>(T + 1) + 2 * padded<uint32_t>(T + 1) + padded<double>(T + 1) + padded<uint32_t>(T + 1) +          // keys, tuples
                         2 * padded<uint32_t>(T + 1) + padded<double>(T + 1) +                                                                   // heads
                         padded<uint32_t>(T + 1) + padded<double>(T + 1) + 5 * padded<uint32_t>(T + 1) + padded<double>(T + 1) +                 // entries, groups
                         padded<double>(R) + 2 * padded<uint32_t>(R + 1) + 2 * padded<uint64_t>(R + 1) +                                          // per read
                         padded<uint64_t>(R + 1) + padded<double>(T + 1) + padded<uint64_t>(T + 2) + padded<uint32_t>(T + 1) +                    // outputs
                         ((sort_bytes + 255) & ~size_t(255)) + ((scan_bytes + 255) & ~size_t(255)) + ((scan64_bytes + 255) & ~size_t(255)) + padded<uint32_t>(1) + 8192;

    DevicePool pool;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&pool.base), bytes, ctx->pool, s));
    const DeviceAllocationGuard allocation_guard{s, pool.base};
    pool.size = bytes;

    uint64_t * d_read_offset = pool.take<uint64_t>(C + 1);
    uint64_t * d_path_offset = pool.take<uint64_t>(C + 1);
    double * d_eff = pool.take<double>(P);
    uint32_t * d_path_group = pool.take<uint32_t>(P);
    uint32_t * d_path_rank = pool.take<uint32_t>(P);
    uint32_t * d_rank_path = pool.take<uint32_t>(P);
    uint32_t * d_source_count = pool.take<uint32_t>(P);
    uint8_t * d_mapq = pool.take<uint8_t>(R);
    uint64_t * d_align_ptr = pool.take<uint64_t>(R + 1);
    int32_t * d_score = pool.take<int32_t>(A);
    uint32_t * d_align_len = pool.take<uint32_t>(A);
    double * d_frag = pool.take<double>(A);
    uint64_t * d_id_ptr = pool.take<uint64_t>(A + 1);
    uint32_t * d_ids = pool.take<uint32_t>(T + 1);
    double * d_phred = pool.take<double>(256);
    uint64_t * d_key_in = pool.take<uint64_t>(T + 1);
    uint64_t * d_key = pool.take<uint64_t>(T + 1);
    uint32_t * d_tuple_in = pool.take<uint32_t>(T + 1);
    uint32_t * d_tuple = pool.take<uint32_t>(T + 1);
    double * d_tuple_lp = pool.take<double>(T + 1);
    uint32_t * d_tuple_len = pool.take<uint32_t>(T + 1);
    uint32_t * d_head = pool.take<uint32_t>(T + 1);
    uint32_t * d_head_pos = pool.take<uint32_t>(T + 1);
    double * d_head_lp = pool.take<double>(T + 1);
    uint32_t * d_entry_rank = pool.take<uint32_t>(T + 1);
    double * d_entry_lp = pool.take<double>(T + 1);
    uint32_t * d_member = pool.take<uint32_t>(T + 1);
    uint32_t * d_group_size = pool.take<uint32_t>(T + 1);
    uint32_t * d_group_start = pool.take<uint32_t>(T + 1);
    uint32_t * d_group_order = pool.take<uint32_t>(T + 1);
    uint32_t * d_staged_ids = pool.take<uint32_t>(T + 1);
    double * d_group_first = pool.take<double>(T + 1);
    double * d_noise = pool.take<double>(R);
    uint32_t * d_n_groups = pool.take<uint32_t>(R + 1);
    uint32_t * d_n_ids = pool.take<uint32_t>(R + 1);
    uint64_t * d_group_base = pool.take<uint64_t>(R + 1);
    uint64_t * d_id_base = pool.take<uint64_t>(R + 1);
    uint64_t * d_out_group_ptr = pool.take<uint64_t>(R + 1);
    double * d_out_group_prob = pool.take<double>(T + 1);
    uint64_t * d_out_id_ptr = pool.take<uint64_t>(T + 2);
    uint32_t * d_out_ids = pool.take<uint32_t>(T + 1);
    char * d_sort_temp = pool.take<char>(sort_bytes);
    char * d_scan_temp = pool.take<char>(scan_bytes);
    char * d_scan64_temp = pool.take<char>(scan64_bytes);
    uint32_t * d_error = pool.take<uint32_t>(1);

    cudaError_t err = (pool.used > pool.size) ? cudaErrorMemoryAllocation : cudaSuccess;

    auto up = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0 && src) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); }
    };

    up(d_read_offset, in->read_offset, (C + 1) * sizeof(uint64_t));
    up(d_path_offset, in->path_offset, (C + 1) * sizeof(uint64_t));
    up(d_eff, in->effective_length, P * sizeof(double));
    up(d_path_group, in->path_group, P * sizeof(uint32_t));
    up(d_path_rank, collapse ? path_rank.data() : nullptr, P * sizeof(uint32_t));
    up(d_rank_path, collapse ? rank_path.data() : nullptr, P * sizeof(uint32_t));
    up(d_source_count, in->source_count, P * sizeof(uint32_t));
    up(d_mapq, in->min_mapq, R * sizeof(uint8_t));
    up(d_align_ptr, in->align_ptr, (R + 1) * sizeof(uint64_t));
    up(d_score, in->score_sum, A * sizeof(int32_t));
    up(d_align_len, in->align_length, A * sizeof(uint32_t));
    up(d_frag, in->frag_log_prob, A * sizeof(double));
    up(d_id_ptr, in->id_ptr, (A + 1) * sizeof(uint64_t));
    up(d_ids, in->ids, T * sizeof(uint32_t));
    up(d_phred, phred_prob, sizeof(phred_prob));

    if (err == cudaSuccess) { err = cudaMemsetAsync(d_error, 0, sizeof(uint32_t), s); }

    uint32_t dev_error = 0;
    uint64_t total_groups = 0, total_ids = 0;

    if (err == cudaSuccess) {

        Alignments al;
        al.n_clusters = C;
        al.n_reads = R;
        al.n_align = A;
        al.n_ids = T;
        al.read_offset = d_read_offset;
        al.path_offset = d_path_offset;
        al.effective_length = d_eff;
        al.path_group = collapse ? d_path_group : nullptr;
        al.path_rank = collapse ? d_path_rank : nullptr;
        al.source_count = in->source_count ? d_source_count : nullptr;
        al.min_mapq = d_mapq;
        al.align_ptr = d_align_ptr;
        al.score_sum = d_score;
        al.align_length = d_align_len;
        al.frag_log_prob = in->frag_log_prob ? d_frag : nullptr;
        al.id_ptr = d_id_ptr;
        al.ids = d_ids;

        const unsigned grid_t = static_cast<unsigned>((T + 1 + 255) / 256);
        const unsigned grid_r = static_cast<unsigned>((R + 1 + 127) / 128);

        if (T > 0) {

            pathTupleKernel<<<grid_t, 256, 0, s>>>(al, d_key_in, d_tuple_in, d_tuple_lp, d_tuple_len, d_error);
            RPVG_LAUNCHED(ctx);

            size_t temp = sort_bytes;
            err = cub::DeviceRadixSort::SortPairs(d_sort_temp, temp, d_key_in, d_key, d_tuple_in, d_tuple, static_cast<int>(T), 0, 32 + read_bits, s);
            ctx->launches += 2 + static_cast<uint64_t>((32 + read_bits + 7) / 8);
        }

        if (err == cudaSuccess) {

            pathHeadKernel<<<grid_t, 256, 0, s>>>(T, d_key, d_tuple, d_tuple_lp, d_tuple_len, d_head, d_head_lp);
            RPVG_LAUNCHED(ctx);

            size_t temp = scan_bytes;
            err = cub::DeviceScan::ExclusiveSum(d_scan_temp, temp, d_head, d_head_pos, static_cast<int>(T + 1), s);
            ctx->launches += 2;
        }

        if (err == cudaSuccess && T > 0) {
            pathCompactKernel<<<grid_t, 256, 0, s>>>(T, d_key, d_head, d_head_pos, d_head_lp, d_entry_rank, d_entry_lp);
            RPVG_LAUNCHED(ctx);
        }

        if (err == cudaSuccess) {

            pathReadKernel<<<grid_r, 128, 0, s>>>(al, min_noise_prob, prob_precision, d_phred, d_head_pos, collapse ? d_rank_path : nullptr, d_entry_rank, d_entry_lp, d_member,
                                                  d_group_first, d_group_size, d_group_start, d_group_order, d_staged_ids, d_noise, d_n_groups, d_n_ids, d_error);
            RPVG_LAUNCHED(ctx);

            size_t temp = scan64_bytes;
            err = cudaMemsetAsync(d_n_groups + R, 0, sizeof(uint32_t), s);
            if (err == cudaSuccess) { err = cudaMemsetAsync(d_n_ids + R, 0, sizeof(uint32_t), s); }
            if (err == cudaSuccess) { err = cub::DeviceScan::ExclusiveSum(d_scan64_temp, temp, d_n_groups, d_group_base, static_cast<int>(R + 1), s); }
            temp = scan64_bytes;
            if (err == cudaSuccess) { err = cub::DeviceScan::ExclusiveSum(d_scan64_temp, temp, d_n_ids, d_id_base, static_cast<int>(R + 1), s); }
            ctx->launches += 4;
        }

        if (err == cudaSuccess) {

            pathEmitKernel<<<grid_r, 128, 0, s>>>(al, d_head_pos, d_group_first, d_group_size, d_group_start, d_group_order, d_staged_ids, d_n_groups, d_group_base, d_id_base,
                                                  d_out_group_ptr, d_out_group_prob, d_out_id_ptr, d_out_ids);
            RPVG_LAUNCHED(ctx);
        }

        if (err == cudaSuccess) { err = cudaMemcpyAsync(&total_groups, d_group_base + R, sizeof(uint64_t), cudaMemcpyDeviceToHost, s); }
        if (err == cudaSuccess) { err = cudaMemcpyAsync(&total_ids, d_id_base + R, sizeof(uint64_t), cudaMemcpyDeviceToHost, s); }
        if (err == cudaSuccess) { err = cudaMemcpyAsync(&dev_error, d_error, sizeof(uint32_t), cudaMemcpyDeviceToHost, s); }
        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }
    }

    auto down = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, s); }
    };

    if (dev_error == 0) {
        down(noise, d_noise, R * sizeof(double));
        down(group_ptr, d_out_group_ptr, (R + 1) * sizeof(uint64_t));
        down(group_prob, d_out_group_prob, total_groups * sizeof(double));
        down(group_id_ptr, d_out_id_ptr, (total_groups + 1) * sizeof(uint64_t));
        down(group_ids, d_out_ids, total_ids * sizeof(uint32_t));
        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }
    }


    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_add_path_probs: ") + cudaGetErrorString(err)); }

    if (dev_error) {
        std::string what = "rpvg_b200_add_path_probs: malformed input:";
        if (dev_error & PATHPROBS_ERR_PATH_RANGE) { what += " path index out of range;"; }
        if (dev_error & PATHPROBS_ERR_NO_PATH) { what += " a read with a noise score below zero has no path of positive effective length (rpvg asserts);"; }
        if (dev_error & PATHPROBS_ERR_LAYOUT) { what += " a read needs at least one alignment path before its noise alignment path, which lists no paths;"; }
        return rpvgFail(ctx, RPVG_B200_ERR_INVALID, what);
    }

    return RPVG_B200_OK;
}
