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
// Values: pairTileKernel. The pairs of all clusters are cut into tiles of 32 x 32 (first group x second group, upper
// triangle of tiles only) and, when the tiles alone would not fill the GPU, into row ranges as well; one 256-thread CTA per
// (tile, row range). The CTA stages the 64 columns of its tile 32 read classes at a time in shared memory (coalesced
// column reads, 0.5 B of L2 traffic per log) and every thread owns 4 pairs: one column read serves 32 pairs, and the
// kernel is bound by arithmetic, not by memory. A log per term is what the reference spends
// (G (G + 1) / 2 x M of them); here sum_r count_r log(x_r) is taken as log(prod_r x_r^count_r): every x_r is split into
// exponent and mantissa (integer operations), the exponents are added as integers (exact), the mantissas in [1, 2)
// multiplied into a running product that is renormalised every 32 read classes, and ONE log is taken per pair and row
// range at the end: ll = ln2 * (sum of exponents) + log(product). The product of n mantissas carries a relative error of
// at most n x 1.1e-16 (1.8e-12 for 16,384 read classes, typically its square root), i.e. an absolute error on ll below
// what one rounding of a sum of that size costs (ulp(1e5) = 1.5e-11). Read classes whose count is not an integer in
// 1..31, and values that are zero, denormal or not finite, take the plain count * log(x) route (a row-uniform branch).
// Deterministic.
// Scan: boundedScanKernel, one CTA per cluster: exclusive prefix maximum in pair order, test, compaction.

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

constexpr uint32_t kPairTile = 32;        // pairs per tile side
constexpr uint32_t kPairRowChunk = 32;    // read classes staged at a time
constexpr int kPairThreads = 128;         // 16 x 8 threads, each 2 first groups x 4 second groups
constexpr uint32_t kPairMaxSplitRows = 16384;   // read classes per work item at most: the exponent sums are 32-bit (31 x 2046 x 16384 < 2^31)

struct PairTileArgs {

    BoundedArgs b;
    const uint4 * items;               // work items: x = cluster, y = tile row | tile column << 16, z = row range, w = row ranges of the tile;
                                       // the row ranges of a tile are adjacent items
    uint32_t rows_per_split;           // read classes per row range (multiple of kPairRowChunk, at most kPairMaxSplitRows)
    double * partial;                  // [n_items x 1024] when rows are split, else NULL
    unsigned long long * log_count;    // terms evaluated (pairs x read classes: for the roofline)
};

// this thread's pairs: first groups i0 and i0 + 16, second groups j0, j0 + 8, j0 + 16, j0 + 24 (positions in `order`);
// pair (a, q) of thread t sits at t * 8 + a * 4 + q of a tile's partial sums
__device__ __forceinline__ void pairsOfThread(const uint32_t tid, const uint32_t ti, const uint32_t tj, uint32_t * i0, uint32_t * j0) {
    *i0 = ti * kPairTile + (tid >> 3);
    *j0 = tj * kPairTile + (tid & 7);
}

__global__ void __launch_bounds__(kPairThreads) pairTileKernel(const PairTileArgs args) {

    // [read class][column of the tile]: 32 first, then 32 second groups. A thread's four second groups lie 8 apart, so the
    // eight threads that share its first groups read consecutive words: no bank conflicts (pitch 65).
    __shared__ double col_s[kPairRowChunk][2 * kPairTile + 1];
    __shared__ double noise_s[kPairRowChunk];
    __shared__ double count_s[kPairRowChunk];
    __shared__ uint32_t power_s[kPairRowChunk];
    __shared__ uint32_t odd_s;

    const BoundedArgs & b = args.b;

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;

    const uint64_t item = blockIdx.x;
    const uint4 what = args.items[item];
    const uint32_t c = what.x;

    const uint64_t g0 = b.group_offset[c];
    const uint32_t G = static_cast<uint32_t>(b.group_offset[c + 1] - g0);
    const uint64_t r0 = b.row_offset[c];
    const uint32_t rows = b.n_classes[c];
    const double * matrix = b.matrix + b.matrix_offset[c];
    const double * noise = b.noise + r0;
    const double * counts = b.counts + r0;
    const uint32_t * order = b.order + g0;

    const uint32_t ti = what.y & 0xffffu, tj = what.y >> 16;
    const uint32_t split = what.z;

    const uint32_t row_begin = split * args.rows_per_split;
    const uint32_t row_end = min(rows, row_begin + args.rows_per_split);

    const uint32_t li = tid >> 3;
    const uint32_t lj = tid & 7;
    uint32_t i0, j0;
    pairsOfThread(tid, ti, tj, &i0, &j0);

    // per pair: product of mantissas, sum of (biased) exponents, and the sum of the terms taken by log
    double product[2][4];
    int32_t exponent[2][4];
    double slow[2][4];

    #pragma unroll
    for (uint32_t a = 0; a < 2; ++a) {
        #pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
            product[a][q] = 1.0;
            exponent[a][q] = 0;
            slow[a][q] = 0;
        }
    }

    // one term: x = m 2^e -> the exponent to the sum, the mantissa (in [1, 2)) into the product
    auto take = [](const double prob, double * prod, int32_t * exp_sum) {
        const int32_t hi = __double2hiint(prob);
        *exp_sum += hi >> 20;
        const double mantissa = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(prob));
        *prod *= mantissa;
        return mantissa;
    };

    if (tid == 0) { odd_s = 0; }
    __syncthreads();

    for (uint32_t base = row_begin; base < row_end; base += kPairRowChunk) {

        const uint32_t n_chunk = min(kPairRowChunk, row_end - base);

        // Kind of read class `lane`, and its place in the staged chunk: first the ones whose count is 1, then counts 2..31,
        // then the ones that take a log — counts that are not small integers, and read classes whose noise is not a
        // normal positive number (with it, noise + a / 2 + b / 2 is one as well for finite a, b >= 0: the fast loops need
        // no test per term). Every warp works this out for itself.
        const bool live = lane < n_chunk;
        const double cnt = live ? counts[base + lane] : 0.0;
        const double nz = live ? noise[base + lane] : 1.0;
        const bool fast = live && cnt >= 1.0 && cnt <= 31.0 && cnt == floor(cnt) && nz >= DBL_MIN && nz <= DBL_MAX;
        const bool once = fast && cnt == 1.0;

        const uint32_t m_once = __ballot_sync(0xffffffffu, once);
        const uint32_t m_more = __ballot_sync(0xffffffffu, fast && !once);
        const uint32_t m_slow = __ballot_sync(0xffffffffu, live && !fast);
        const uint32_t below = (1u << lane) - 1u;
        const uint32_t n_once = __popc(m_once), n_more = __popc(m_more);

        uint32_t slot = n_chunk + (lane - n_chunk);   // (lanes past the chunk keep their place)
        if (once) { slot = __popc(m_once & below); }
        else if (fast) { slot = n_once + __popc(m_more & below); }
        else if (live) { slot = n_once + n_more + __popc(m_slow & below); }

        const uint32_t power_total = __reduce_add_sync(0xffffffffu, fast ? static_cast<uint32_t>(cnt) : 0u);

        __syncthreads();   // the previous chunk has been consumed

        // stage: warp w loads columns w, w + 4, ..: lane = read class (coalesced)
        for (uint32_t k = warp; k < 2 * kPairTile; k += kPairThreads / 32) {
            const uint32_t pos = (k < kPairTile ? ti * kPairTile + k : tj * kPairTile + (k - kPairTile));
            double v = 0;
            if (pos < G && live) { v = matrix[static_cast<uint64_t>(order[pos]) * rows + base + lane]; }
            col_s[slot][k] = v;
            // (anything but a finite value >= 0: from here on this CTA takes the log of every term, like the reference)
            if (!(v >= 0.0 && v <= DBL_MAX)) { odd_s = 1; }
        }
        if (warp == 0) {
            noise_s[slot] = nz;
            count_s[slot] = cnt;
            power_s[slot] = fast ? static_cast<uint32_t>(cnt) : 0u;
        }

        __syncthreads();

        const bool odd = odd_s != 0;
        const uint32_t n1 = odd ? 0u : n_once;
        const uint32_t n2 = odd ? 0u : n_once + n_more;

        // read classes seen once
        #pragma unroll 2
        for (uint32_t r = 0; r < n1; ++r) {

            // noise + col_first / 2 + col_second / 2, in the reference's order
            const double nzr = noise_s[r];
            const double first0 = nzr + col_s[r][li] / 2.0;
            const double first1 = nzr + col_s[r][li + 16] / 2.0;

            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
                const double second = col_s[r][kPairTile + lj + 8 * q];
                take(first0 + second / 2.0, &product[0][q], &exponent[0][q]);
                take(first1 + second / 2.0, &product[1][q], &exponent[1][q]);
            }
        }

        // read classes seen 2..31 times: x^count from the squares of the mantissa
        for (uint32_t r = n1; r < n2; ++r) {

            const uint32_t power = power_s[r];
            const double nzr = noise_s[r];
            const double first0 = nzr + col_s[r][li] / 2.0;
            const double first1 = nzr + col_s[r][li + 16] / 2.0;

            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {

                const double second = col_s[r][kPairTile + lj + 8 * q];

                #pragma unroll
                for (uint32_t a = 0; a < 2; ++a) {

                    const double prob = (a == 0 ? first0 : first1) + second / 2.0;
                    const int32_t hi = __double2hiint(prob);
                    exponent[a][q] += static_cast<int32_t>(power) * (hi >> 20);

                    const double mantissa = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(prob));
                    double square = mantissa * mantissa;
                    double factor = (power & 1u) ? mantissa : 1.0;
                    if (power & 2u) { factor *= square; }

                    if (power >= 4) {
                        square *= square;
                        if (power & 4u) { factor *= square; }
                        if (power >= 8) {
                            square *= square;
                            if (power & 8u) { factor *= square; }
                            if (power >= 16) {
                                square *= square;
                                factor *= square;
                            }
                        }
                    }

                    product[a][q] *= factor;
                }
            }
        }

        // the terms that take a log
        for (uint32_t r = n2; r < n_chunk; ++r) {

            const double weight = count_s[r];
            const double nzr = noise_s[r];
            const double first0 = nzr + col_s[r][li] / 2.0;
            const double first1 = nzr + col_s[r][li + 16] / 2.0;

            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
                const double second = col_s[r][kPairTile + lj + 8 * q];
                slow[0][q] = fma(weight, log(first0 + second / 2.0), slow[0][q]);
                slow[1][q] = fma(weight, log(first1 + second / 2.0), slow[1][q]);
            }
        }

        // renormalise (the products grew by at most 2^(31 x 32)) and take the bias of the chunk's exponents off
        const int32_t bias = odd ? 0 : static_cast<int32_t>(1023u * power_total);

        #pragma unroll
        for (uint32_t a = 0; a < 2; ++a) {
            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
                const int32_t hi = __double2hiint(product[a][q]);
                exponent[a][q] += (hi >> 20) - 1023 - bias;
                product[a][q] = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(product[a][q]));
            }
        }
    }

    // ln2 in two pieces (the high one has 21 zero bits at the end: its product with the exponent sum is exact up to 2^21)
    double acc[2][4];
    #pragma unroll
    for (uint32_t a = 0; a < 2; ++a) {
        #pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
            const double e = static_cast<double>(exponent[a][q]);
            acc[a][q] = fma(e, 6.93147180369123816490e-01, fma(e, 1.90821492927058770002e-10, log(product[a][q]) + slow[a][q]));
        }
    }

    if (tid == 0 && args.log_count) {
        const uint32_t ni = min(kPairTile, G - ti * kPairTile), nj = min(kPairTile, G - tj * kPairTile);
        const unsigned long long pairs = (ti == tj) ? static_cast<unsigned long long>(ni) * (ni + 1) / 2 : static_cast<unsigned long long>(ni) * nj;
        atomicAdd(args.log_count, pairs * (row_end > row_begin ? row_end - row_begin : 0));
    }

    if (args.partial) {

        double * out = args.partial + item * (kPairTile * kPairTile) + tid * 8;
        #pragma unroll
        for (uint32_t a = 0; a < 2; ++a) {
            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) { out[a * 4 + q] = acc[a][q]; }
        }
        return;
    }

    const uint64_t p0 = b.pair_offset[c];

    #pragma unroll
    for (uint32_t a = 0; a < 2; ++a) {

        const uint32_t i = i0 + 16 * a;
        if (i >= G) { continue; }

        const uint32_t ga = order[i];
        const uint64_t row_start = static_cast<uint64_t>(i) * G - (static_cast<uint64_t>(i) * (i - 1)) / 2;

        #pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
            const uint32_t j = j0 + 8 * q;
            if (j < G && j >= i) {
                const uint32_t gb = order[j];
                const uint64_t p = p0 + row_start + (j - i);
                b.cand_loglik[p] = acc[a][q];
                b.pair_ll[p] = acc[a][q] + (b.log_freq[g0 + ga] + b.log_freq[g0 + gb] + (ga == gb ? 0.0 : 0.6931471805599453));
            }
        }
    }
}

// Row ranges were summed apart: add them in order, per pair, and write the values (one CTA per tile: launched over all
// items, only the first row range of a tile works).
__global__ void __launch_bounds__(kPairThreads) pairCombineKernel(const PairTileArgs args) {

    const BoundedArgs & b = args.b;

    const uint64_t item = blockIdx.x;
    const uint4 what = args.items[item];
    if (what.z != 0) { return; }

    const uint32_t c = what.x;
    const uint32_t n_splits = what.w;

    const uint64_t g0 = b.group_offset[c];
    const uint32_t G = static_cast<uint32_t>(b.group_offset[c + 1] - g0);
    const uint32_t * order = b.order + g0;

    const uint32_t ti = what.y & 0xffffu, tj = what.y >> 16;

    const uint32_t tid = threadIdx.x;
    uint32_t i0, j0;
    pairsOfThread(tid, ti, tj, &i0, &j0);

    const uint64_t p0 = b.pair_offset[c];

    for (uint32_t a = 0; a < 2; ++a) {

        const uint32_t i = i0 + 16 * a;
        if (i >= G) { continue; }

        const uint32_t ga = order[i];
        const uint64_t row_start = static_cast<uint64_t>(i) * G - (static_cast<uint64_t>(i) * (i - 1)) / 2;

        for (uint32_t q = 0; q < 4; ++q) {

            const uint32_t j = j0 + 8 * q;
            if (j >= G || j < i) { continue; }

            double acc = 0, comp = 0;
            for (uint32_t sp = 0; sp < n_splits; ++sp) {
                const double y = args.partial[(item + sp) * (kPairTile * kPairTile) + tid * 8 + a * 4 + q] - comp;
                const double t = acc + y;
                comp = (t - acc) - y;
                acc = t;
            }

            const uint32_t gb = order[j];
            const uint64_t p = p0 + row_start + (j - i);
            b.cand_loglik[p] = acc;
            b.pair_ll[p] = acc + (b.log_freq[g0 + ga] + b.log_freq[g0 + gb] + (ga == gb ? 0.0 : 0.6931471805599453));
        }
    }
}

// ---- the reference's bound on a whole row of pairs (rpvg:src/path_estimator.cpp:414, 427-433) ---------------------------
// max over the columns of every read class, halved (read_path_probs.rowwise().maxCoeff() / group_size): one thread per
// read class, coalesced over the read classes of a column
__global__ void pairRowMaxKernel(const BoundedArgs b, const uint32_t n_clusters, const uint64_t n_rows_total, double * __restrict__ row_max) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_rows_total) { return; }

    const uint32_t c = clusterOf(b.row_offset, n_clusters, r);
    const uint32_t local = static_cast<uint32_t>(r - b.row_offset[c]);
    const uint32_t rows = b.n_classes[c];
    if (local >= rows) { return; }

    const uint32_t G = static_cast<uint32_t>(b.group_offset[c + 1] - b.group_offset[c]);
    const double * matrix = b.matrix + b.matrix_offset[c];

    double best = matrix[local];
    for (uint32_t g = 1; g < G; ++g) { best = fmax(best, matrix[static_cast<uint64_t>(g) * rows + local]); }

    row_max[r] = best / 2.0;
}

// optimal log-likelihood of any pair whose first group is order[i]: one warp per (cluster, i)
__global__ void __launch_bounds__(kBoundedThreads) pairRowBoundKernel(const BoundedArgs b, const uint32_t n_clusters, const uint64_t n_groups_total,
                                                                      const double * __restrict__ row_max, double * __restrict__ bound) {

    const uint32_t lane = threadIdx.x & 31;
    const uint64_t k = (blockIdx.x * static_cast<uint64_t>(kBoundedThreads) + threadIdx.x) >> 5;
    if (k >= n_groups_total) { return; }

    const uint32_t c = clusterOf(b.group_offset, n_clusters, k);
    const uint64_t g0 = b.group_offset[c];
    const uint64_t r0 = b.row_offset[c];
    const uint32_t rows = b.n_classes[c];
    const uint32_t a = b.order[k];

    const double * col = b.matrix + b.matrix_offset[c] + static_cast<uint64_t>(a) * rows;

    double acc = 0;
    for (uint32_t r = lane; r < rows; r += 32) {
        const double base = b.noise[r0 + r] + col[r] / 2.0;
        acc += b.counts[r0 + r] * log(base + row_max[r0 + r]);
    }
    acc = warpSum(acc);

    if (lane == 0) { bound[k] = acc + (b.log_freq[g0 + a] + 0.6931471805599453); }
}

__global__ void pairFillKernel(const uint64_t n, double * __restrict__ pair_ll) {

    const uint64_t p = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (p < n) { pair_ll[p] = -DBL_MAX; }
}

// largest value among the pairs of the first `head_rows` rows of every cluster (a prefix of its pairs): one CTA per cluster
__global__ void __launch_bounds__(kBoundedThreads) pairHeadMaxKernel(const BoundedArgs b, const uint32_t head_rows, double * __restrict__ head_max) {

    __shared__ double red[kBoundedThreads / 32];

    const uint32_t c = blockIdx.x;
    const uint64_t G = b.group_offset[c + 1] - b.group_offset[c];
    const uint64_t h = G < head_rows ? G : head_rows;
    const uint64_t n = h * G - (h * (h - 1)) / 2;
    const double * pair_ll = b.pair_ll + b.pair_offset[c];

    double best = -DBL_MAX;
    for (uint64_t p = threadIdx.x; p < n; p += kBoundedThreads) { best = fmax(best, pair_ll[p]); }

    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) { best = fmax(best, __shfl_xor_sync(0xffffffffu, best, off)); }
    if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = best; }
    __syncthreads();

    if (threadIdx.x == 0) {
        for (int w = 1; w < kBoundedThreads / 32; ++w) { best = fmax(best, red[w]); }
        head_max[c] = best;
    }
}

constexpr int kScanThreads = 1024;
constexpr uint32_t kScanItems = 4;   // consecutive pairs per thread and round

// One CTA per cluster: exclusive prefix maximum in pair order, the test against it, compaction of the pairs that pass.
// A round takes 4096 pairs (four consecutive ones per thread: a cluster of 5,000 groups has 12.5 M of them).
__global__ void __launch_bounds__(kScanThreads) boundedScanKernel(const BoundedArgs args) {

    __shared__ double warp_max[kScanThreads / 32];
    __shared__ uint32_t warp_count[kScanThreads / 32];
    __shared__ double carry;
    __shared__ uint32_t kept_total;

    const uint32_t c = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    constexpr uint32_t num_warps = kScanThreads / 32;

    const uint64_t p0 = args.pair_offset[c];
    const uint64_t n_pairs = args.pair_offset[c + 1] - p0;

    double * pair_ll = args.pair_ll + p0;
    double * raw = args.cand_loglik + p0;   // raw sums of all pairs first, compacted in place
    uint32_t * cand_pair = args.cand_pair + p0;

    if (tid == 0) {
        carry = -DBL_MAX;
        kept_total = 0;
    }
    __syncthreads();

    for (uint64_t base = 0; base < n_pairs; base += kScanThreads * kScanItems) {

        const uint64_t first = base + static_cast<uint64_t>(tid) * kScanItems;

        double ll[kScanItems], raw_value[kScanItems];
        #pragma unroll
        for (uint32_t i = 0; i < kScanItems; ++i) {
            const uint64_t p = first + i;
            ll[i] = p < n_pairs ? pair_ll[p] : -DBL_MAX;
            raw_value[i] = p < n_pairs ? raw[p] : 0.0;
        }

        // maximum of this thread's pairs, exclusive scan of the maxima over the warp and the warps before it
        double mine = ll[0];
        #pragma unroll
        for (uint32_t i = 1; i < kScanItems; ++i) { mine = fmax(mine, ll[i]); }

        double incl = mine;
        #pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const double up = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) { incl = fmax(incl, up); }
        }
        double before = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) { before = -DBL_MAX; }

        if (lane == 31) { warp_max[warp] = incl; }
        __syncthreads();

        before = fmax(before, carry);
        for (uint32_t w = 0; w < warp; ++w) { before = fmax(before, warp_max[w]); }

        // the test, pair by pair, against everything before the pair
        bool keep[kScanItems];
        uint32_t n_keep = 0;
        #pragma unroll
        for (uint32_t i = 0; i < kScanItems; ++i) {
            keep[i] = (first + i < n_pairs) && !(ll[i] - before < args.min_log_diff);
            n_keep += keep[i] ? 1 : 0;
            before = fmax(before, ll[i]);
        }

        uint32_t count_incl = n_keep;
        #pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, count_incl, off);
            if (lane >= off) { count_incl += up; }
        }

        if (lane == 31) { warp_count[warp] = count_incl; }
        __syncthreads();   // (also: all reads of raw[] / carry / kept_total of this round are done)

        uint32_t offset = kept_total + count_incl - n_keep;
        for (uint32_t w = 0; w < warp; ++w) { offset += warp_count[w]; }

        __syncthreads();

        #pragma unroll
        for (uint32_t i = 0; i < kScanItems; ++i) {
            if (keep[i]) {
                cand_pair[offset] = static_cast<uint32_t>(first + i);
                raw[offset] = raw_value[i];   // offset <= pair index: never overwrites a value a later round still needs
                ++offset;
            }
        }

        if (tid == kScanThreads - 1) {
            double round_max = carry;
            for (uint32_t w = 0; w < num_warps; ++w) { round_max = fmax(round_max, warp_max[w]); }
            carry = round_max;
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
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

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

        // ---- plan. The reference skips all pairs of a first group whose bound (its log-likelihood with the best column of
        // every read class as partner) is already out of reach of the running maximum. The running maximum is the prefix
        // maximum over the pairs in enumeration order, so any maximum over an initial stretch of rows is a lower bound
        // of it for every later row: rows whose bound fails against it (with the margin) are skipped by the reference
        // as well. Two waves: the first 32 groups of every cluster (the most probable ones: the caller orders them by
        // marginal posterior), then only the tile rows that still hold a row in reach. Pairs never evaluated stay at the
        // most negative value and drop out in the scan.
        std::vector<uint32_t> n_classes(C, 0);
        err = cudaMemcpyAsync(n_classes.data(), grouped->d_n_classes, 4 * static_cast<size_t>(C), cudaMemcpyDeviceToHost, s);
        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

        const uint64_t R_total = grouped->h_rows_total;

        char * extra = nullptr;
        const size_t extra_bytes = pad(8 * R_total) + pad(8 * G) + pad(8 * C) + 256;
        if (err == cudaSuccess) { err = cudaMallocFromPoolAsync(reinterpret_cast<void **>(&extra), extra_bytes, ctx->pool, s); }

        double * d_row_max = reinterpret_cast<double *>(extra);
        double * d_bound = reinterpret_cast<double *>(extra + pad(8 * R_total));
        double * d_head_max = reinterpret_cast<double *>(extra + pad(8 * R_total) + pad(8 * G));
        unsigned long long * d_log_count = reinterpret_cast<unsigned long long *>(extra + extra_bytes - 256);

        cudaEvent_t ev0 = nullptr, ev1 = nullptr;
        if (err == cudaSuccess) { err = cudaEventCreate(&ev0); }
        if (err == cudaSuccess) { err = cudaEventCreate(&ev1); }
        float pair_ms = 0;

        std::vector<double> bound(G, 0), head_max(C, 0);

        // one wave: the tiles in `tiles` (cluster, tile row, tile column), cut into row ranges while they leave SMs idle
        auto run_wave = [&](const std::vector<uint4> & tiles) {

            if (tiles.empty() || err != cudaSuccess) { return; }

            uint32_t most_rows = 0;
            for (auto & t: tiles) { most_rows = std::max(most_rows, n_classes[t.x]); }

            const uint64_t want_items = 4ull * static_cast<uint64_t>(ctx->num_sms);
            uint32_t rows_per_split = std::max(most_rows, kPairRowChunk);

            if (tiles.size() < want_items) {
                const uint64_t splits = (want_items + tiles.size() - 1) / tiles.size();
                rows_per_split = static_cast<uint32_t>((most_rows + splits - 1) / splits);
                rows_per_split = std::max(8 * kPairRowChunk, ((rows_per_split + kPairRowChunk - 1) / kPairRowChunk) * kPairRowChunk);
            }

            rows_per_split = std::min(rows_per_split, kPairMaxSplitRows);
            rows_per_split = ((rows_per_split + kPairRowChunk - 1) / kPairRowChunk) * kPairRowChunk;

            std::vector<uint4> items;
            items.reserve(tiles.size());
            bool split_rows = false;

            for (auto & t: tiles) {
                const uint32_t splits = std::max(1u, (n_classes[t.x] + rows_per_split - 1) / rows_per_split);
                split_rows = split_rows || splits > 1;
                for (uint32_t sp = 0; sp < splits; ++sp) { items.push_back(make_uint4(t.x, t.y, sp, splits)); }
            }

            if (items.size() >= (1ull << 31)) { err = cudaErrorInvalidValue; return; }

            char * wave = nullptr;
            const size_t wave_bytes = pad(16 * items.size()) + (split_rows ? pad(8 * items.size() * kPairTile * kPairTile) : 0);
            err = cudaMallocFromPoolAsync(reinterpret_cast<void **>(&wave), wave_bytes, ctx->pool, s);
            if (err != cudaSuccess) { return; }

            uint4 * d_items = reinterpret_cast<uint4 *>(wave);
            err = cudaMemcpyAsync(d_items, items.data(), 16 * items.size(), cudaMemcpyHostToDevice, s);

            PairTileArgs tile_args;
            tile_args.b = args;
            tile_args.items = d_items;
            tile_args.rows_per_split = rows_per_split;
            tile_args.partial = split_rows ? reinterpret_cast<double *>(wave + pad(16 * items.size())) : nullptr;
            tile_args.log_count = d_log_count;

            if (err == cudaSuccess) {

                cudaEventRecord(ev0, s);
                pairTileKernel<<<static_cast<unsigned>(items.size()), kPairThreads, 0, s>>>(tile_args);
                ctx->launches++;
                if (split_rows) {
                    pairCombineKernel<<<static_cast<unsigned>(items.size()), kPairThreads, 0, s>>>(tile_args);
                    ctx->launches++;
                }
                cudaEventRecord(ev1, s);
                err = cudaGetLastError();
            }

            // (the item list is a host vector: done with it before it goes)
            if (err == cudaSuccess) { err = cudaEventSynchronize(ev1); }
            if (err == cudaSuccess) {
                float ms = 0;
                cudaEventElapsedTime(&ms, ev0, ev1);
                pair_ms += ms;
            }

            cudaFreeAsync(wave, s);
        };

        if (err == cudaSuccess) {

            err = cudaMemsetAsync(d_log_count, 0, 8, s);

            if (err == cudaSuccess && R_total > 0) {
                pairRowMaxKernel<<<static_cast<unsigned>((R_total + 255) / 256), 256, 0, s>>>(args, C, R_total, d_row_max);
                pairRowBoundKernel<<<static_cast<unsigned>((G * 32 + kBoundedThreads - 1) / kBoundedThreads), kBoundedThreads, 0, s>>>(args, C, G, d_row_max, d_bound);
                pairFillKernel<<<static_cast<unsigned>((NP + 255) / 256), 256, 0, s>>>(NP, d_pair_ll);
                ctx->launches += 3;
                err = cudaGetLastError();
            }

            // wave 1: tile row 0 of every cluster
            std::vector<uint4> tiles;
            for (uint32_t c = 0; c < C; ++c) {
                const uint32_t side = static_cast<uint32_t>((group_offset[c + 1] - group_offset[c] + kPairTile - 1) / kPairTile);
                for (uint32_t tj = 0; tj < side; ++tj) { tiles.push_back(make_uint4(c, tj << 16, 0, 1)); }
            }
            run_wave(tiles);

            if (err == cudaSuccess) {
                pairHeadMaxKernel<<<C, kBoundedThreads, 0, s>>>(args, kPairTile, d_head_max);
                ctx->launches++;
                err = cudaGetLastError();
            }
            if (err == cudaSuccess) { err = cudaMemcpyAsync(bound.data(), d_bound, 8 * G, cudaMemcpyDeviceToHost, s); }
            if (err == cudaSuccess) { err = cudaMemcpyAsync(head_max.data(), d_head_max, 8 * static_cast<size_t>(C), cudaMemcpyDeviceToHost, s); }
            if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

            // wave 2: the other tile rows that hold a first group still in reach
            tiles.clear();
            for (uint32_t c = 0; c < C && err == cudaSuccess; ++c) {
                const uint64_t g0 = group_offset[c];
                const uint32_t Gc = static_cast<uint32_t>(group_offset[c + 1] - g0);
                const uint32_t side = (Gc + kPairTile - 1) / kPairTile;
                for (uint32_t ti = 1; ti < side; ++ti) {
                    bool in_reach = false;
                    for (uint32_t i = ti * kPairTile; i < std::min(Gc, (ti + 1) * kPairTile) && !in_reach; ++i) {
                        in_reach = !(bound[g0 + i] - head_max[c] < min_log_diff);
                    }
                    if (!in_reach) { continue; }
                    for (uint32_t tj = ti; tj < side; ++tj) { tiles.push_back(make_uint4(c, ti | (tj << 16), 0, 1)); }
                }
            }
            run_wave(tiles);

            if (err == cudaSuccess) {
                boundedScanKernel<<<C, kScanThreads, 0, s>>>(args);
                ctx->launches++;
                err = cudaGetLastError();
            }

            unsigned long long log_count = 0;
            if (err == cudaSuccess) { err = cudaMemcpyAsync(&log_count, d_log_count, 8, cudaMemcpyDeviceToHost, s); }
            if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

            if (err == cudaSuccess) {
                grouped->pair_ms += pair_ms;
                grouped->pair_logs += static_cast<double>(log_count);
            }
        }

        if (ev0) { cudaEventDestroy(ev0); }
        if (ev1) { cudaEventDestroy(ev1); }
        if (extra) { cudaFreeAsync(extra, s); }
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


extern "C" int rpvg_b200_grouped_pair_stats(const rpvg_b200_grouped * grouped, double * kernel_ms, double * log_evaluations) {

    if (!grouped) { return RPVG_B200_ERR_INVALID; }
    if (kernel_ms) { *kernel_ms = grouped->pair_ms; }
    if (log_evaluations) { *log_evaluations = grouped->pair_logs; }
    return RPVG_B200_OK;
}
