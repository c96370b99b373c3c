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
// (G (G + 1) / 2 x M of them); here sum_r count_r log(x_r) is taken as log(prod_r x_r^count_r). Every read class has a
// power of two s_r = 2^-k_r that brings all its terms below 1 (from the reference's own row maximum, which the bound
// kernels need anyway): the scaled terms are multiplied into a running product per pair (scaling by a power of two is
// exact), sum_r count_r k_r is the same integer for all pairs, and the products are renormalised (exponent moved to an
// integer sum) whenever the smallest value they could have reached nears the end of the double range. ONE log per pair
// and row range at the end: ll = ln2 * (sum of exponents) + log(product). A term costs two fp64 instructions instead of
// the ~50 of a log. The product of n factors carries a relative error of at most n x 1.1e-16 (1.8e-12 for 16,384 read
// classes, typically its square root), i.e. an absolute error on ll below what one rounding of a sum of that size
// costs (ulp(1e5) = 1.5e-11). Read classes whose count is not an integer in 1..31, whose values span too many binary
// orders or are negative or not finite take the count * log(x) route (a row-uniform branch). The columns of the next 32
// read classes are fetched by asynchronous copies while a chunk is worked on. Deterministic.
// Scan: boundedScanKernel, one CTA per cluster: exclusive prefix maximum in pair order, test, compaction.

#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
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
    const double * row_max;            // per read class: largest column / 2 (pairRowMaxKernel)
    unsigned long long * log_count;    // terms evaluated (pairs x read classes: for the roofline)
};

// this thread's pairs: first groups i0 and i0 + 16, second groups j0, j0 + 8, j0 + 16, j0 + 24 (positions in `order`);
// pair (a, q) of thread t sits at t * 8 + a * 4 + q of a tile's partial sums
__device__ __forceinline__ void pairsOfThread(const uint32_t tid, const uint32_t ti, const uint32_t tj, uint32_t * i0, uint32_t * j0) {
    *i0 = ti * kPairTile + (tid >> 3);
    *j0 = tj * kPairTile + (tid & 7);
}

// log(x) for the terms that keep their log (counts that are not small integers): x = m 2^k with m in [1/sqrt2, sqrt2),
// log m = 2 atanh(s), s = (m - 1) / (m + 1), |s| <= 0.1716, by its power series up to s^21 (the next term is below 1e-17);
// the quotient from the hardware's reciprocal seed and two Newton steps plus a correction of the quotient itself. About
// 30 instructions against the ~100 issue slots of the library's log, to 1-2 ulp of the result. Anything but a normal
// positive number goes to the library.
__device__ __forceinline__ double pairLog(const double x) {

    const int32_t hi = __double2hiint(x);
    if (static_cast<uint32_t>(hi - 0x00100000) >= 0x7fe00000u) { return log(x); }

    // mantissa in [1, 2); above sqrt2 it is halved and the exponent goes up by one
    int32_t k = (hi >> 20) - 1023;
    int32_t mhi = (hi & 0x000fffff) | 0x3ff00000;
    if (mhi >= 0x3ff6a09f) {   // high word of sqrt2, rounded up
        mhi -= 0x00100000;
        k += 1;
    }
    const double m = __hiloint2double(mhi, __double2loint(x));

    const double f = m - 1.0;
    const double d = m + 1.0;

    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = fma(r, fma(-d, r, 1.0), r);
    r = fma(r, fma(-d, r, 1.0), r);

    double q = f * r;
    q = fma(r, fma(-d, q, f), q);   // q = f / d to the last bit or two

    const double z = q * q;
    double p = 2.0 / 21.0;
    p = fma(p, z, 2.0 / 19.0);
    p = fma(p, z, 2.0 / 17.0);
    p = fma(p, z, 2.0 / 15.0);
    p = fma(p, z, 2.0 / 13.0);
    p = fma(p, z, 2.0 / 11.0);
    p = fma(p, z, 2.0 / 9.0);
    p = fma(p, z, 2.0 / 7.0);
    p = fma(p, z, 2.0 / 5.0);
    p = fma(p, z, 2.0 / 3.0);

    // 2 q + q z p, with what the rounding of q lost put back: f - q d is exact enough in one fma
    const double tail = fma(q * z, p, 2.0 * r * fma(-d, q, f));
    const double kd = static_cast<double>(k);
    return fma(kd, 6.93147180369123816490e-01, fma(kd, 1.90821492927058770002e-10, tail + 2.0 * q) );
}

// t^P for the small counts that are most of the repeated read classes
template <uint32_t P>
__device__ __forceinline__ double pairPower(const double t) {
    if (P == 2) { return t * t; }
    if (P == 3) { return t * t * t; }
    const double t2 = t * t;
    if (P == 4) { return t2 * t2; }
    if (P == 5) { return t2 * t2 * t; }
    return t2 * t2 * t2;   // 6
}

// One staged chunk: 32 read classes of the tile's 64 columns, and what the loops need to know about each read class
struct PairStage {
    // [read class][column of the tile]: 32 first, then 32 second groups. A thread's four second groups lie 8 apart, so the
    // eight threads that share its first groups read consecutive words: no bank conflicts (pitch 65).
    double col[kPairRowChunk][2 * kPairTile + 1];
    double2 noise_half[kPairRowChunk];   // x: noise of the read class, times its scale on the fast routes; y: scale / 2 (1 / 2
                                         // for the read classes that take logs) — one 16-byte load
    double count[kPairRowChunk];
    int32_t floor_[kPairRowChunk];    // count x binary orders the scaled terms of the read class can lie below 1
    uint32_t power[kPairRowChunk];
};

// 8 bytes global -> shared without a register in between
__device__ __forceinline__ void pairCopy8(void * dst, const void * src) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(d), "l"(src) : "memory");
}

__global__ void __launch_bounds__(kPairThreads) pairTileKernel(const PairTileArgs args) {

    // two stages: the columns of the next 32 read classes are on their way (asynchronous copies) while this chunk is
    // worked on
    __shared__ PairStage stages[2];
    __shared__ uint64_t column_s[2 * kPairTile];   // offset of every column of the tile in the matrix; ~0: past the last group

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
    const double * row_max = args.row_max + r0;
    const uint32_t * order = b.order + g0;

    const uint32_t ti = what.y & 0xffffu, tj = what.y >> 16;
    const uint32_t split = what.z;

    const uint32_t row_begin = split * args.rows_per_split;
    const uint32_t row_end = min(rows, row_begin + args.rows_per_split);

    const uint32_t li = tid >> 3;
    const uint32_t lj = tid & 7;
    uint32_t i0, j0;
    pairsOfThread(tid, ti, tj, &i0, &j0);

    if (tid < 2 * kPairTile) {
        const uint32_t pos = (tid < kPairTile ? ti * kPairTile + tid : tj * kPairTile + (tid - kPairTile));
        column_s[tid] = pos < G ? static_cast<uint64_t>(order[pos]) * rows : ~0ull;
    }

    // per pair: the running product, the exponents taken out of it, and the sum of the terms taken by log; for all pairs
    // alike: sum of count x k over the read classes of the fast routes
    double product[2][4];
    int32_t exponent[2][4];
    double slow[2][4];
    double common_exponent = 0;

    #pragma unroll
    for (uint32_t a = 0; a < 2; ++a) {
        #pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
            product[a][q] = 1.0;
            exponent[a][q] = 0;
            slow[a][q] = 0;
        }
    }

    // binary orders the products may have lost since they were last brought back to [1, 2)
    int32_t depth = 0;
    constexpr int32_t kDepthLimit = 900;

    auto renormalise = [&]() {
        #pragma unroll
        for (uint32_t a = 0; a < 2; ++a) {
            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
                const int32_t hi = __double2hiint(product[a][q]);
                exponent[a][q] += (hi >> 20) - 1023;
                product[a][q] = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(product[a][q]));
            }
        }
        depth = 0;
    };

    // what a warp knows about read class `lane` of a chunk before it sends for its columns
    struct Meta { double count, noise, top; };

    auto load_meta = [&](const uint32_t base) {
        Meta m = {0.0, 1.0, 1.0};
        if (base < row_end && base + lane < row_end) {
            m.count = counts[base + lane];
            m.noise = noise[base + lane];
            m.top = m.noise + 2.0 * row_max[base + lane];   // (NaN where the read class holds a value that is not finite and >= 0)
        }
        return m;
    };

    // the chunk's split into the three routes, and sum of count x k over its fast read classes
    struct Plan { uint32_t n_once, n_more; double shift; };

    // Sends for the columns of the chunk at `base` into stage `buffer`; the read classes are placed by route: first the
    // ones whose count is 1, then counts 2..31, then the ones that take a log. A fast read class has noise and
    // noise + 2 x (largest column / 2) normal positive numbers: all its terms lie between the two. Scale 2^-k brings the
    // upper one into [1/2, 1); the lower one then sits `orders` binary orders below 1 at most. Every warp works the
    // placement out for itself.
    auto send = [&](const uint32_t base, const uint32_t buffer, const Meta & m) {

        PairStage & st = stages[buffer];
        const uint32_t n_chunk = min(kPairRowChunk, row_end - base);
        const bool live = lane < n_chunk;

        const int32_t top_order = (__double2hiint(m.top) >> 20) - 1023;
        const int32_t low_order = (__double2hiint(m.noise) >> 20) - 1023;
        const int32_t orders = top_order + 1 - low_order;

        const bool fast = live && m.count >= 1.0 && m.count <= 31.0 && m.count == floor(m.count) && m.noise >= DBL_MIN && m.top <= 1e300 && m.top >= m.noise &&
                          orders * static_cast<int32_t>(m.count) <= kDepthLimit / 4;
        const bool once = fast && m.count == 1.0;

        const uint32_t m_once = __ballot_sync(0xffffffffu, once);
        const uint32_t m_more = __ballot_sync(0xffffffffu, fast && !once);
        const uint32_t m_slow = __ballot_sync(0xffffffffu, live && !fast);
        const uint32_t below = (1u << lane) - 1u;

        Plan plan;
        plan.n_once = __popc(m_once);
        plan.n_more = __popc(m_more);

        uint32_t slot = lane;   // (lanes past the chunk keep their place)
        if (once) { slot = __popc(m_once & below); }
        else if (fast) { slot = plan.n_once + __popc(m_more & below); }
        else if (live) { slot = plan.n_once + plan.n_more + __popc(m_slow & below); }

        // (sum over the fast read classes of the chunk, the same in every lane: count x k are small integers, exact)
        double shift = fast ? m.count * static_cast<double>(top_order + 1) : 0.0;
        #pragma unroll
        for (int off = 16; off > 0; off >>= 1) { shift += __shfl_xor_sync(0xffffffffu, shift, off); }
        plan.shift = shift;

        // warp w sends for columns w, w + 4, ..: lane = read class (coalesced)
        for (uint32_t k = warp; k < 2 * kPairTile; k += kPairThreads / 32) {
            const uint64_t column = column_s[k];
            if (live && column != ~0ull) { pairCopy8(&st.col[slot][k], matrix + column + base + lane); }
            else { st.col[slot][k] = 0.0; }
        }

        if (warp == 0) {
            // scale = 2^-(top_order + 1)
            const double scale = fast ? __hiloint2double((1023 - (top_order + 1)) << 20, 0) : 1.0;
            st.noise_half[slot] = make_double2(m.noise * scale, 0.5 * scale);
            st.count[slot] = m.count;
            st.floor_[slot] = fast ? orders * static_cast<int32_t>(m.count) : 0;
            st.power[slot] = fast ? static_cast<uint32_t>(m.count) : 0u;
        }

        asm volatile("cp.async.commit_group;" ::: "memory");
        return plan;
    };

    __syncthreads();   // column_s

    Plan plan = send(row_begin, 0, load_meta(row_begin));
    Meta ahead = load_meta(row_begin + kPairRowChunk);

    uint32_t buffer = 0;

    for (uint32_t base = row_begin; base < row_end; base += kPairRowChunk, buffer ^= 1u) {

        const uint32_t n_chunk = min(kPairRowChunk, row_end - base);

        // the next chunk into the other stage (its last readers passed the barrier at the end of the previous round)
        Plan next_plan = {0, 0, 0.0};
        if (base + kPairRowChunk < row_end) {
            next_plan = send(base + kPairRowChunk, buffer ^ 1u, ahead);
            ahead = load_meta(base + 2 * kPairRowChunk);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }

        __syncthreads();   // this chunk is in, for everybody

        const PairStage & st = stages[buffer];
        const uint32_t n1 = plan.n_once;
        const uint32_t n2 = plan.n_once + plan.n_more;
        common_exponent += plan.shift;

        // term of a pair = (noise + first / 2) + second / 2 in the reference's order, scaled by the read class's power of two
        // (exactly): fma(second, scale / 2, fma(first, scale / 2, noise x scale))

        // read classes seen once, four at a time (each is at most a quarter of the depth limit deep)
        auto once_row = [&](const uint32_t r) {

            const double2 nh = st.noise_half[r];
            const double hs = nh.y;
            const double nzr = nh.x;
            const double first0 = fma(st.col[r][li], hs, nzr);
            const double first1 = fma(st.col[r][li + 16], hs, nzr);

            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
                const double second = st.col[r][kPairTile + lj + 8 * q];
                product[0][q] *= fma(second, hs, first0);
                product[1][q] *= fma(second, hs, first1);
            }
        };

        uint32_t r = 0;

        for (; r + 4 <= n1; r += 4) {

            const int32_t deeper = (st.floor_[r] + st.floor_[r + 1]) + (st.floor_[r + 2] + st.floor_[r + 3]);
            if (depth + deeper > kDepthLimit) { renormalise(); }
            depth += deeper;

            once_row(r);
            once_row(r + 1);
            once_row(r + 2);
            once_row(r + 3);
        }

        for (; r < n1; ++r) {
            if (depth + st.floor_[r] > kDepthLimit) { renormalise(); }
            depth += st.floor_[r];
            once_row(r);
        }

        // read classes seen 2..31 times: x^count, by a product for the smallest counts, from the squares otherwise
        for (; r < n2; ++r) {

            if (depth + st.floor_[r] > kDepthLimit) { renormalise(); }
            depth += st.floor_[r];

            const uint32_t power = st.power[r];
            const double2 nh = st.noise_half[r];
            const double hs = nh.y;
            const double nzr = nh.x;
            const double first0 = fma(st.col[r][li], hs, nzr);
            const double first1 = fma(st.col[r][li + 16], hs, nzr);

            auto each_term = [&](auto raise) {
                #pragma unroll
                for (uint32_t q = 0; q < 4; ++q) {
                    const double second = st.col[r][kPairTile + lj + 8 * q];
                    product[0][q] *= raise(fma(second, hs, first0));
                    product[1][q] *= raise(fma(second, hs, first1));
                }
            };

            switch (power) {
                case 2: each_term([](const double t) { return pairPower<2>(t); }); break;
                case 3: each_term([](const double t) { return pairPower<3>(t); }); break;
                case 4: each_term([](const double t) { return pairPower<4>(t); }); break;
                case 5: each_term([](const double t) { return pairPower<5>(t); }); break;
                case 6: each_term([](const double t) { return pairPower<6>(t); }); break;
                default:
                    each_term([power](const double t) {
                        double square = t * t;
                        double factor = (power & 1u) ? t : 1.0;
                        if (power & 2u) { factor *= square; }
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
                        return factor;
                    });
            }
        }

        // the terms that take a log (unscaled: half = 1 / 2)
        for (; r < n_chunk; ++r) {

            const double weight = st.count[r];
            const double nzr = st.noise_half[r].x;
            const double first0 = nzr + st.col[r][li] / 2.0;
            const double first1 = nzr + st.col[r][li + 16] / 2.0;

            #pragma unroll
            for (uint32_t q = 0; q < 4; ++q) {
                const double second = st.col[r][kPairTile + lj + 8 * q];
                slow[0][q] = fma(weight, pairLog(first0 + second / 2.0), slow[0][q]);
                slow[1][q] = fma(weight, pairLog(first1 + second / 2.0), slow[1][q]);
            }
        }

        plan = next_plan;

        __syncthreads();   // the stage may be refilled
    }

    renormalise();

    // ln2 in two pieces (the high one has 21 zero bits at the end: its product with the exponent sum is exact up to 2^21)
    double acc[2][4];
    #pragma unroll
    for (uint32_t a = 0; a < 2; ++a) {
        #pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
            const double e = common_exponent + static_cast<double>(exponent[a][q]);
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

                const double part = args.partial[(item + sp) * (kPairTile * kPairTile) + tid * 8 + a * 4 + q];

                // (a row range with a zero term is -inf: plain addition, the compensation would turn it into NaN)
                if (!(fabs(part) <= DBL_MAX) || !(fabs(acc) <= DBL_MAX)) {
                    acc += part;
                    comp = 0;
                    continue;
                }

                const double y = part - comp;
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

    // (a read class that holds anything but finite values >= 0 gets NaN: the pair kernel then takes the plain log of its
    // terms, like the reference, and the bound of every row of pairs it is in compares false: nothing is ruled out)
    double best = 0;
    bool plain = true;
    for (uint32_t g = 0; g < G; ++g) {
        const double v = matrix[static_cast<uint64_t>(g) * rows + local];
        plain = plain && v >= 0.0 && v <= DBL_MAX;
        best = fmax(best, v);
    }

    row_max[r] = plain ? best / 2.0 : __longlong_as_double(0x7ff8000000000000ll);
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

    // stage times on stderr when RPVG_B200_BATCH_TRACE is set (each lap synchronises the stream: tracing only)
    const bool trace = std::getenv("RPVG_B200_BATCH_TRACE") != nullptr;
    auto lap_from = std::chrono::steady_clock::now();
    auto lap = [&](const char * what) {
        if (!trace) { return; }
        cudaStreamSynchronize(s);
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[rpvg_b200 pairs] %s: %.3f ms\n", what, std::chrono::duration<double, std::milli>(now - lap_from).count());
        lap_from = now;
    };

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

            const uint64_t want_items = 32ull * static_cast<uint64_t>(ctx->num_sms);
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
            tile_args.row_max = d_row_max;

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

            lap("row maxima, bounds, fill");

            // wave 1: tile row 0 of every cluster
            std::vector<uint4> tiles;
            for (uint32_t c = 0; c < C; ++c) {
                const uint32_t side = static_cast<uint32_t>((group_offset[c + 1] - group_offset[c] + kPairTile - 1) / kPairTile);
                for (uint32_t tj = 0; tj < side; ++tj) { tiles.push_back(make_uint4(c, tj << 16, 0, 1)); }
            }
            run_wave(tiles);
            lap("wave 1");

            if (err == cudaSuccess) {
                pairHeadMaxKernel<<<C, kBoundedThreads, 0, s>>>(args, kPairTile, d_head_max);
                ctx->launches++;
                err = cudaGetLastError();
            }
            if (err == cudaSuccess) { err = cudaMemcpyAsync(bound.data(), d_bound, 8 * G, cudaMemcpyDeviceToHost, s); }
            if (err == cudaSuccess) { err = cudaMemcpyAsync(head_max.data(), d_head_max, 8 * static_cast<size_t>(C), cudaMemcpyDeviceToHost, s); }
            if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

            lap("head maxima to the host");

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
            lap("wave 2");

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

    lap("scan");

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

    lap("candidates to the host");

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
