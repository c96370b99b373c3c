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
//   2. greedy scan, one thread per problem (sequential by definition: the predicate is against the running
//      representative, which is not an equivalence relation);
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

// One thread per problem. sorted[] holds the row ids of every problem in sorted order (problems contiguous).
// Writes, per sorted position: class index within its problem; per problem: number of classes.
static __global__ void collapseScanKernel(const CollapseRows rows, const uint32_t * __restrict__ sorted, const double prob_precision,
                                   uint32_t * __restrict__ sorted_class, uint32_t * __restrict__ n_classes) {

    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= rows.n_problems) { return; }

    const uint64_t s0 = rows.problem_row_offset[p];
    const uint64_t s1 = rows.problem_row_offset[p + 1];

    uint32_t classes = 0;
    uint32_t rep = 0;

    for (uint64_t s = s0; s < s1; ++s) {

        const uint32_t row = sorted[s];

        if (s == s0 || !collapseWithinPrecision(rows, rep, row, prob_precision)) {
            rep = row;
            classes++;
        }
        sorted_class[s] = classes - 1;
    }

    n_classes[p] = classes;
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
