// Read-class collapse on the device (shared by collapse.cu and the subset-EM pipeline).
//
// Reference semantics restated (never its code): rowSortProbabilityMatrix + probabilityCountRowSorter
// (rpvg:src/path_estimator.cpp:13-31,197-217): rows ordered lexicographically over all columns (noise column last), two
// values being "equal" under Utils::doubleCompare (relative 100 eps, rpvg:src/utils.hpp:81-93), ties broken by count;
// then readCollapseProbabilityMatrix (:219-259): a greedy scan in which a row joins the CURRENT representative iff every
// column differs by less than prob_precision, counts add up, the representative keeps its values.
//
// Rows are sparse here (CSR, problem-local column indices ascending; an absent column is the dense matrix's 0), the noise
// column is a separate per-row value compared after the path columns.
//   1. per problem, a bitonic sorting network over the row ids with the reference's comparator (one CTA per problem).
//      A network is data-oblivious: the tolerance comparator is not always a strict weak ordering (a ~ b, b ~ c, a < c),
//      and a merge-path sort fed such a comparator can partition out of bounds (seen with cub::DeviceMergeSort); a
//      network always yields a permutation. Ties under the comparator are broken by row id, so the result is the one a
//      stable sort would give whenever the comparator is consistent;
//   2. greedy scan (sequential by definition: the predicate is against the running representative, which is not an
//      equivalence relation), cut into independent stretches at the rows that certainly start a class;
//   3. class ids / representatives / class counts.
// What is exact: the partition into classes, the class counts and the class order. Rows the comparator calls equal
// (all columns within 100 eps, same count) are ordered arbitrarily by the reference's unstable std::sort; here they keep
// input order, so a representative may be a different member of such a tie (values equal within 100 eps).
#pragma once

#include <cfloat>

#include "common.cuh"

struct CollapseRows {

    uint32_t n_problems;
    uint64_t n_rows;
    const uint64_t * problem_row_offset;   // [n_problems + 1]
    const uint32_t * row_problem;          // [n_rows]
    const uint64_t * row_ptr;              // [n_rows + 1]
    const uint32_t * col;
    const double * val;
    const double * noise;                  // [n_rows]
    const double * cnt;                    // [n_rows]
};

__device__ __forceinline__ bool collapseDoubleCompare(const double a, const double b) {

#ifdef RPVG_B200_COLLAPSE_EXACT_COMPARE
    return a == b;   // debugging aid only
#else
    return (a == b) || (fabs(a - b) < fabs(fmin(a, b)) * (DBL_EPSILON * 100));
#endif
}

struct CollapseRowLess {

    CollapseRows rows;

    __device__ bool operator()(const uint32_t a, const uint32_t b) const {

        uint64_t ea = rows.row_ptr[a], eb = rows.row_ptr[b];
        const uint64_t ea1 = rows.row_ptr[a + 1], eb1 = rows.row_ptr[b + 1];

        while (ea < ea1 || eb < eb1) {

            const uint32_t ca = ea < ea1 ? rows.col[ea] : 0xffffffffu;
            const uint32_t cb = eb < eb1 ? rows.col[eb] : 0xffffffffu;

            double va = 0, vb = 0;
            if (ca <= cb) { va = rows.val[ea++]; }
            if (cb <= ca) { vb = rows.val[eb++]; }

            if (!collapseDoubleCompare(va, vb)) { return va < vb; }
        }

        if (!collapseDoubleCompare(rows.noise[a], rows.noise[b])) { return rows.noise[a] < rows.noise[b]; }
        if (!collapseDoubleCompare(rows.cnt[a], rows.cnt[b])) { return rows.cnt[a] < rows.cnt[b]; }
        return false;
    }
};

constexpr int kCollapseSortThreads = 512;

// Bitonic network over keys[0, n) run by one CTA of kCollapseSortThreads threads (all compare-exchanges ascending, so the
// virtual +inf padding above n stays put). `less` is the reference's comparator on two keys; keys the comparator calls
// equal are ordered by key value.
template <typename Less>
__device__ __forceinline__ void collapseBitonicNetwork(uint32_t * keys, const uint32_t n, const Less & less) {

    if (n < 2) { return; }

    uint32_t padded = 2;
    while (padded < n) { padded <<= 1; }
    const uint32_t half = padded >> 1;

    auto exchange = [&](const uint32_t lo, const uint32_t hi) {
        if (hi < n) {
            const uint32_t a = keys[lo];
            const uint32_t b = keys[hi];
            // b before a: strictly less, or tied under the comparator with the smaller key
            if (less(b, a) || (!less(a, b) && b < a)) {
                keys[lo] = b;
                keys[hi] = a;
            }
        }
    };

    for (uint32_t k = 2; k <= padded; k <<= 1) {

        const uint32_t kh = k >> 1;
        for (uint32_t i = threadIdx.x; i < half; i += kCollapseSortThreads) {
            const uint32_t block = i / kh, off = i % kh;
            exchange(block * k + off, block * k + k - 1 - off);
        }
        __syncthreads();

        for (uint32_t j = k >> 2; j >= 1; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < half; i += kCollapseSortThreads) {
                const uint32_t lo = (i / j) * 2 * j + (i % j);
                exchange(lo, lo + j);
            }
            __syncthreads();
        }
    }
}

// One CTA per problem: sorted[s0, s1) holds the problem's row ids on entry, in the reference's order on exit.
static __global__ void __launch_bounds__(kCollapseSortThreads) collapseSortKernel(const CollapseRows rows, uint32_t * __restrict__ sorted) {

    const uint32_t p = blockIdx.x;
    const uint64_t s0 = rows.problem_row_offset[p];
    const uint32_t n = static_cast<uint32_t>(rows.problem_row_offset[p + 1] - s0);

    CollapseRowLess less;
    less.rows = rows;

    collapseBitonicNetwork(sorted + s0, n, less);
}

// true iff every column (absent = 0) and the noise value of rows a and b differ by less than prob_precision
__device__ __forceinline__ bool collapseWithinPrecision(const CollapseRows & rows, const uint32_t a, const uint32_t b, const double prob_precision) {

    uint64_t ea = rows.row_ptr[a], eb = rows.row_ptr[b];
    const uint64_t ea1 = rows.row_ptr[a + 1], eb1 = rows.row_ptr[b + 1];

    while (ea < ea1 || eb < eb1) {

        const uint32_t ca = ea < ea1 ? rows.col[ea] : 0xffffffffu;
        const uint32_t cb = eb < eb1 ? rows.col[eb] : 0xffffffffu;

        double va = 0, vb = 0;
        if (ca <= cb) { va = rows.val[ea++]; }
        if (cb <= ca) { vb = rows.val[eb++]; }

        if (fabs(va - vb) >= prob_precision) { return false; }
    }

    return fabs(rows.noise[a] - rows.noise[b]) < prob_precision;
}

// The greedy scan, in three parallel passes. sorted[] holds the row ids of every problem in sorted order (problems
// contiguous); sorted_class[] receives, per sorted position, the class index within its problem; n_classes per problem.
//
// A row joins the CURRENT representative iff all columns are within the precision of the representative's — sequential by
// definition. But the chain is cut wherever a row is certainly not within the precision of any representative it could
// meet: if it differs from its predecessor by at least twice the precision in some column, then (triangle inequality) it
// differs from the predecessor's representative by more than the precision, and from the predecessor itself too — it
// starts a class whatever happened before. Pass 1 marks these certain starts (one thread per adjacent pair of rows);
// pass 2 walks every stretch between two certain starts with the exact sequential rule (a warp per stretch, the lanes
// sharing out the columns of each comparison); pass 3 numbers the classes (one CTA per problem). The result is the sequential scan's, bit for bit.

// true iff some column (absent = 0) or the noise value of rows a and b differ by at least `distance`
__device__ __forceinline__ bool collapseFartherThan(const CollapseRows & rows, const uint32_t a, const uint32_t b, const double distance) {

    uint64_t ea = rows.row_ptr[a], eb = rows.row_ptr[b];
    const uint64_t ea1 = rows.row_ptr[a + 1], eb1 = rows.row_ptr[b + 1];

    while (ea < ea1 || eb < eb1) {

        const uint32_t ca = ea < ea1 ? rows.col[ea] : 0xffffffffu;
        const uint32_t cb = eb < eb1 ? rows.col[eb] : 0xffffffffu;

        double va = 0, vb = 0;
        if (ca <= cb) { va = rows.val[ea++]; }
        if (cb <= ca) { vb = rows.val[eb++]; }

        if (fabs(va - vb) >= distance) { return true; }
    }

    return fabs(rows.noise[a] - rows.noise[b]) >= distance;
}

// pass 1: flag[s] = 1 where a class certainly starts (first row of a problem, or farther than twice the precision, with a
// hair of slack for the rounding of the differences, from the row before), else 0
static __global__ void collapseBoundaryKernel(const CollapseRows rows, const uint32_t * __restrict__ sorted, const double prob_precision, uint32_t * __restrict__ flag) {

    const uint64_t s = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (s >= rows.n_rows) { return; }

    const uint32_t row = sorted[s];
    const uint32_t p = rows.row_problem[row];

    bool start = (s == rows.problem_row_offset[p]);
    if (!start) { start = collapseFartherThan(rows, sorted[s - 1], row, 2.0 * prob_precision * (1.0 + 1e-9)); }

    flag[s] = start ? 1u : 0u;
}

// pass 2: every stretch (from a certain start up to the next one or the end of the problem) is walked with the
// sequential rule; rows that become representatives inside the stretch get flag 2. A warp takes 32 positions, and the
// stretches that begin there and are longer than one row one after the other, all lanes on one comparison: the lanes
// share out the entries of the row, look each column up in the representative (binary search in its sorted columns) and
// vote. A column the representative has and the row lacks differs by the representative's own value: the
// representative's entries that are at least the precision are counted once, and all of them must have found a partner.
// Same predicate as collapseWithinPrecision, entry for entry.
constexpr int kCollapseStretchThreads = 128;

static __global__ void __launch_bounds__(kCollapseStretchThreads) collapseStretchKernel(const CollapseRows rows, const uint32_t * __restrict__ sorted, const double prob_precision,
                                                                                        uint32_t * flag) {

    const uint32_t lane = threadIdx.x & 31;
    const uint64_t first = ((blockIdx.x * static_cast<uint64_t>(kCollapseStretchThreads) + threadIdx.x) >> 5) * 32;
    const uint64_t s = first + lane;

    // (flag[] holds 0 / 1 from pass 1; a 2 is only ever written inside the writer's own stretch)
    bool head = false;
    uint64_t stretch_end = 0;
    if (s < rows.n_rows && flag[s] == 1u) {
        stretch_end = rows.problem_row_offset[rows.row_problem[sorted[s]] + 1];
        head = s + 1 < stretch_end && flag[s + 1] != 1u;
    }

    uint32_t heads = __ballot_sync(0xffffffffu, head);

    while (heads != 0) {

        const int src = __ffs(heads) - 1;
        heads &= heads - 1;

        const uint64_t s0 = first + src;
        const uint64_t s1 = __shfl_sync(0xffffffffu, stretch_end, src);

        uint32_t rep = sorted[s0];
        uint64_t a0 = 0, a1 = 0;
        uint32_t rep_large = 0;
        bool rep_loaded = false;

        for (uint64_t t = s0 + 1; t < s1; ++t) {

            if (flag[t] == 1u) { break; }

            if (!rep_loaded) {
                a0 = rows.row_ptr[rep];
                a1 = rows.row_ptr[rep + 1];
                uint32_t large = 0;
                for (uint64_t e = a0 + lane; e < a1; e += 32) { large += fabs(rows.val[e]) >= prob_precision ? 1u : 0u; }
                rep_large = __reduce_add_sync(0xffffffffu, large);
                rep_loaded = true;
            }

            const uint32_t row = sorted[t];
            const uint64_t b0 = rows.row_ptr[row], b1 = rows.row_ptr[row + 1];

            bool differs = false;
            uint32_t partners = 0;

            for (uint64_t base = b0; base < b1; base += 32) {

                const uint64_t e = base + lane;
                if (e < b1) {

                    const uint32_t column = rows.col[e];
                    uint64_t lo = a0, hi = a1;
                    while (lo < hi) {
                        const uint64_t mid = (lo + hi) >> 1;
                        if (rows.col[mid] < column) { lo = mid + 1; } else { hi = mid; }
                    }

                    double of_rep = 0;
                    if (lo < a1 && rows.col[lo] == column) {
                        of_rep = rows.val[lo];
                        partners += fabs(of_rep) >= prob_precision ? 1u : 0u;
                    }
                    if (fabs(of_rep - rows.val[e]) >= prob_precision) { differs = true; }
                }
            }

            differs = __any_sync(0xffffffffu, differs);
            if (!differs) { differs = __reduce_add_sync(0xffffffffu, partners) != rep_large; }
            if (!differs) { differs = !(fabs(rows.noise[rep] - rows.noise[row]) < prob_precision); }

            if (differs) {
                rep = row;
                rep_loaded = false;
                if (lane == 0) { flag[t] = 2u; }
            }
        }

        __syncwarp();
    }
}

constexpr int kCollapseIndexThreads = 256;

// pass 3: class index = number of starts up to and including the position, minus one (in place), and the class count
static __global__ void __launch_bounds__(kCollapseIndexThreads) collapseIndexKernel(const CollapseRows rows, uint32_t * sorted_class, uint32_t * __restrict__ n_classes) {

    __shared__ uint32_t warp_total[kCollapseIndexThreads / 32];
    __shared__ uint32_t carry;

    const uint32_t p = blockIdx.x;
    const uint64_t s0 = rows.problem_row_offset[p];
    const uint64_t s1 = rows.problem_row_offset[p + 1];

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) { carry = 0; }
    __syncthreads();

    for (uint64_t base = s0; base < s1; base += kCollapseIndexThreads) {

        const uint64_t s = base + tid;
        const uint32_t start = (s < s1 && sorted_class[s] != 0u) ? 1u : 0u;

        uint32_t incl = start;
        #pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) { incl += up; }
        }

        if (lane == 31) { warp_total[warp] = incl; }
        __syncthreads();

        uint32_t before = carry;
        for (uint32_t w = 0; w < warp; ++w) { before += warp_total[w]; }

        if (s < s1) { sorted_class[s] = before + incl - 1; }

        __syncthreads();
        if (tid == kCollapseIndexThreads - 1) { carry = before + incl; }
        __syncthreads();
    }

    if (tid == 0) { n_classes[p] = carry; }
}

// Per sorted position: scatter to per-row class ids, mark representatives, add the counts up per class (in sorted
// order, by the thread of the representative, so the sum order is fixed).
static __global__ void collapseClassesKernel(const CollapseRows rows, const uint32_t * __restrict__ sorted, const uint32_t * __restrict__ sorted_class,
                                      const uint64_t * __restrict__ class_offset, uint32_t * __restrict__ row_class, uint32_t * __restrict__ class_rep,
                                      double * __restrict__ class_count) {

    const uint64_t s = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (s >= rows.n_rows) { return; }

    const uint32_t row = sorted[s];
    const uint32_t p = rows.row_problem[row];
    const uint32_t cls = sorted_class[s];

    if (row_class) { row_class[row] = cls; }

    const bool first = (s == rows.problem_row_offset[p]) || sorted_class[s - 1] != cls;

    if (first) {

        const uint64_t s1 = rows.problem_row_offset[p + 1];
        double total = 0;
        for (uint64_t t = s; t < s1 && sorted_class[t] == cls; ++t) { total += rows.cnt[sorted[t]]; }

        class_rep[class_offset[p] + cls] = row;
        class_count[class_offset[p] + cls] = total;
    }
}
