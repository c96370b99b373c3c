// emStreamKernel — EM of ONE giant path cluster on the whole GPU, the matrix streamed from HBM every iteration.
//
// The sliced-ELL kernels (em_sell.cuh) keep a cluster in the shared memory of at most 16 SMs (one thread-block cluster);
// a path cluster far beyond that (hundreds of thousands of entries and more: deep metagenomic strain sets, a whole
// pantranscriptome left as one component) would crawl there. Here all SMs work on it: a cooperative launch of one
// persistent grid, the read classes and the paths each cut into one contiguous range per CTA of about equal entries,
// two grid-wide barriers per iteration:
//
//   E-step  rows of the CTA's range, 32 read classes per warp at a time: the warp loads their (contiguous) entries
//           together, coalesced, 128 per step, the products go through shared memory and every lane sums its own read
//           class in entry order: w_i = count_i / (sum_j p_ij a_j + noise_i a_noise) into global memory; the CTA's share of
//           sum_i noise_i w_i into partial[cta]
//   grid barrier
//   M-step  columns of the CTA's range, 8 lanes per path over the column-major copy: a_j <- a_j * sum_i p_ij w_i / total
//           in place (the E-step is over); every CTA adds partial[] up in the same order and gets the same noise
//           abundance; the CTA's convergence vote into flags[cta]
//   grid barrier
//   every CTA reads all votes: same count of converged iterations everywhere, same stopping iteration.
//
// Bytes per iteration (what the HBM roofline of this kernel is quoted on): both orientations are read, 12 B per entry each
// (value + index), plus 24 B per read class (noise, count, offset) + 8 B (w written) + 8 B per entry gathered from w (L2),
// plus 24 B per path. DESIGN.md's algorithmic figure counts the matrix once (12 nnz + 20 M + 16 (N + 1)): a deterministic
// scheme needs both orientations, so half of the HBM rate is this kernel's ceiling on that scale. Fixed ranges, fixed
// summation trees, no atomics: bit-reproducible, and the arithmetic is the one of emClusterKernel (the reference's
// operation order: (a * sum) / total, |da| / a > conv).
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"
#include "em_kernels.cuh"

constexpr int kStreamThreads = 512;
constexpr uint32_t kStreamMaxCtas = 2048;
constexpr uint32_t kStreamStep = 128;   // entries a warp multiplies per step of the E-step

struct StreamArgs {

    EmArgs em;
    uint32_t cluster;
    double * partial;        // [grid] noise partial of every CTA
    uint32_t * flags;        // [grid] convergence vote of every CTA
    double * low_partial;    // [grid] mass below min_em_abundance (final pass)
};

__global__ void __launch_bounds__(kStreamThreads, 2) emStreamKernel(const StreamArgs sa) {

    __shared__ double red[32];
    __shared__ double product_area[kStreamThreads / 32][kStreamStep];

    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();

    const EmArgs & args = sa.em;

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    double * products = product_area[warp];
    const uint32_t cta = blockIdx.x;
    const uint32_t n_ctas = gridDim.x;

    if (*args.error) { return; }   // uniform over the grid

    const uint32_t c = sa.cluster;

    const uint64_t r0 = args.read_offset[c];
    const uint32_t M = static_cast<uint32_t>(args.read_offset[c + 1] - r0);
    const uint64_t p0 = args.path_offset[c];
    const uint32_t N = static_cast<uint32_t>(args.path_offset[c + 1] - p0);

    const uint64_t * row_ptr = args.row_ptr + r0;
    const double * noise = args.noise + r0;
    const double * cnt = args.cnt + r0;
    const uint64_t * col_ptr = args.col_ptr + p0;

    double * a = args.abund + p0 + c;   // N + 1 relative abundances, noise last
    double * w = args.w + r0;
    double * abundances = args.abundances + p0;

    // ---- ranges of this CTA: rows and columns, each about 1 / n_ctas of the entries (+ one unit per row / column)
    const uint64_t e0 = row_ptr[0];
    const uint64_t Z = row_ptr[M] - e0;
    const uint64_t k0 = col_ptr[0];

    uint32_t row_lo, row_hi, col_lo, col_hi;
    {
        // (offset + index is strictly increasing: rows and columns without entries are spread as well)
        auto split = [](const uint64_t * ptr, uint32_t n, uint64_t base, uint64_t total, uint32_t part, uint32_t parts) {
            if (part == 0) { return 0u; }
            if (part >= parts) { return n; }
            const uint64_t target = (total * part) / parts;
            uint32_t lo = 0, hi = n;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if ((ptr[mid] - base) + mid >= target) { hi = mid; } else { lo = mid + 1; }
            }
            return lo;
        };
        row_lo = split(row_ptr, M, e0, Z + M, cta, n_ctas);
        row_hi = split(row_ptr, M, e0, Z + M, cta + 1, n_ctas);
        col_lo = split(col_ptr, N, k0, Z + N, cta, n_ctas);
        col_hi = split(col_ptr, N, k0, Z + N, cta + 1, n_ctas);
    }

    // ---- total read count: every CTA sums its own rows, then all add the partial sums in the same order (counts are
    // integers: the sum is exact whatever the order)
    double count_sum = 0;
    for (uint32_t i = row_lo + tid; i < row_hi; i += kStreamThreads) { count_sum += cnt[i]; }
    const double cta_count = blockSum(count_sum, red);
    if (tid == 0) { sa.low_partial[cta] = cta_count; }

    grid.sync();

    double count_part = 0;
    for (uint32_t q = tid; q < n_ctas; q += kStreamThreads) { count_part += sa.low_partial[q]; }
    const double total = blockSum(count_part, red);

    // rpvg:src/path_abundance_estimator.cpp:54 — single-precision quotient widened to double
    const double a_init = static_cast<double>(1 / static_cast<float>(N + 1));
    for (uint32_t j = col_lo + tid; j < col_hi; j += kStreamThreads) { a[j] = a_init; }

    double a_noise = a_init;

    const double conv = args.params.max_rel_em_conv;
    const double min_abundance = args.params.min_em_abundance;

    grid.sync();

    uint32_t conv_its = 0;
    uint32_t it = 0;

    while (it < args.params.max_em_its) {

        // ---- E-step over this CTA's rows
        double noise_acc = 0;

        // A warp takes 32 consecutive read classes at a time (lane = read class). Their entries are one contiguous stretch
        // of the row-major arrays: the lanes load it together, 128 entries a step — fully coalesced, every load independent
        // of the row offsets — multiply by the gathered abundances and park the products in shared memory; then every lane
        // adds up the products of its own read class, in entry order (the order a sequential loop over the row has).
        // (the next tile's offsets, noise and count are loaded while this tile is worked on; the entry arrays are read
        // once per iteration and marked streaming, so that they do not push a and w out of the L2)
        constexpr uint32_t kTileStride = (kStreamThreads / 32) * 32;

        uint32_t tile = row_lo + warp * 32;
        uint64_t begin = 0, end = 0;
        double nz = 0, ct = 0;
        if (tile + lane < row_hi) {
            begin = row_ptr[tile + lane];
            end = row_ptr[tile + lane + 1];
            nz = noise[tile + lane];
            ct = cnt[tile + lane];
        }

        for (; tile < row_hi; tile += kTileStride) {

            const uint32_t i = tile + lane;
            const bool live = i < row_hi;

            const uint32_t next = i + kTileStride;
            uint64_t next_begin = 0, next_end = 0;
            double next_nz = 0, next_ct = 0;
            if (next < row_hi) {
                next_begin = row_ptr[next];
                next_end = row_ptr[next + 1];
                next_nz = noise[next];
                next_ct = cnt[next];
            }

            const uint32_t last = min(31u, row_hi - 1 - tile);
            const uint64_t tile_begin = __shfl_sync(0xffffffffu, begin, 0);
            const uint64_t tile_end = __shfl_sync(0xffffffffu, end, last);

            double s = 0;

            for (uint64_t step = tile_begin; step < tile_end; step += kStreamStep) {

                // all loads of the step first, then all gathers: the lanes past the end repeat the last entry (their
                // products are never added)
                double v[kStreamStep / 32];
                uint32_t x[kStreamStep / 32];

                #pragma unroll
                for (uint32_t u = 0; u < kStreamStep / 32; ++u) {
                    const uint64_t e = min(step + u * 32 + lane, tile_end - 1);
                    v[u] = __ldcs(args.val + e);
                    x[u] = __ldcs(args.col + e);
                }

                double g[kStreamStep / 32];

                #pragma unroll
                for (uint32_t u = 0; u < kStreamStep / 32; ++u) { g[u] = a[x[u]]; }

                #pragma unroll
                for (uint32_t u = 0; u < kStreamStep / 32; ++u) { products[u * 32 + lane] = v[u] * g[u]; }

                __syncwarp();

                const uint64_t from = max(begin, step);
                const uint64_t to = min(end, step + kStreamStep);
                for (uint64_t e = from; e < to; ++e) { s += products[e - step]; }

                __syncwarp();
            }

            if (live) {
                const double wi = ct / (s + nz * a_noise);
                w[i] = wi;
                noise_acc += nz * wi;
            }

            begin = next_begin;
            end = next_end;
            nz = next_nz;
            ct = next_ct;
        }

        const double cta_noise = blockSum(noise_acc, red);
        if (tid == 0) { sa.partial[cta] = cta_noise; }

        grid.sync();

        // ---- noise column: the partials of all CTAs, added in the same order everywhere
        double part = 0;
        for (uint32_t q = tid; q < n_ctas; q += kStreamThreads) { part += sa.partial[q]; }
        const double noise_g = blockSum(part, red);

        const double a_noise_new = (a_noise * noise_g) / total;
        bool ok = !(a_noise_new >= min_abundance && fabs(a_noise_new - a_noise) / a_noise_new > conv);

        // ---- M-step over this CTA's columns: 8 lanes per path, or a whole warp where the range holds few, long columns
        // (with 8 lanes the last round of a short range would leave most groups idle)
        if ((col_hi - col_lo) * 8 < 4 * kStreamThreads) {

            for (uint32_t j = col_lo + warp; j < col_hi; j += kStreamThreads / 32) {

                const uint64_t k1 = col_ptr[j + 1];
                double gsum = 0;
                #pragma unroll 4
                for (uint64_t k = col_ptr[j] + lane; k < k1; k += 32) { gsum += __ldcs(args.csc_val + k) * w[__ldcs(args.csc_row + k)]; }
                gsum = warpSum(gsum);

                if (lane == 0) {
                    const double a_old = a[j];
                    const double a_new = (a_old * gsum) / total;
                    a[j] = a_new;
                    if (a_new >= min_abundance && fabs(a_new - a_old) / a_new > conv) { ok = false; }
                }
            }

        } else {

            const uint32_t sub = tid & 7;

            // (the loop bound is the warp's: its four columns may end at different places, the shuffles want all lanes)
            for (uint32_t jw = col_lo + warp * 4; jw < col_hi; jw += kStreamThreads / 8) {

                const uint32_t j = jw + (lane >> 3);
                const bool live = j < col_hi;

                double gsum = 0;
                if (live) {
                    const uint64_t k1 = col_ptr[j + 1];
                    #pragma unroll 4
                    for (uint64_t k = col_ptr[j] + sub; k < k1; k += 8) { gsum += __ldcs(args.csc_val + k) * w[__ldcs(args.csc_row + k)]; }
                }

                gsum += __shfl_xor_sync(0xffffffffu, gsum, 4);
                gsum += __shfl_xor_sync(0xffffffffu, gsum, 2);
                gsum += __shfl_xor_sync(0xffffffffu, gsum, 1);

                if (live && sub == 0) {
                    const double a_old = a[j];
                    const double a_new = (a_old * gsum) / total;
                    a[j] = a_new;
                    if (a_new >= min_abundance && fabs(a_new - a_old) / a_new > conv) { ok = false; }
                }
            }
        }

        const int cta_ok = __syncthreads_and(ok);
        if (tid == 0) { sa.flags[cta] = cta_ok ? 1u : 0u; }

        a_noise = a_noise_new;
        ++it;

        grid.sync();

        // ---- the grid's verdict: all votes, read by everyone
        int all = 1;
        for (uint32_t q = tid; q < n_ctas; q += kStreamThreads) { all &= sa.flags[q] != 0; }

        if (__syncthreads_and(all)) {
            if (++conv_its == args.params.min_em_conv_its) { break; }
        } else {
            conv_its = 0;
        }
    }

    // ---- rpvg:src/path_abundance_estimator.cpp:100-113, every CTA on its own columns
    double noise_part = 0;
    for (uint32_t j = col_lo + tid; j < col_hi; j += kStreamThreads) {

        const double aj = a[j];
        if (aj < min_abundance) {
            noise_part += aj * total;
            abundances[j] = 0;
        } else {
            abundances[j] = aj * total;
        }
    }

    const double low_sum = blockSum(noise_part, red);
    if (tid == 0) { sa.low_partial[cta] = low_sum; }

    grid.sync();

    if (cta == 0) {

        double low = 0;
        for (uint32_t q = tid; q < n_ctas; q += kStreamThreads) { low += sa.low_partial[q]; }
        const double low_total = blockSum(low, red);

        if (tid == 0) {
            a[N] = a_noise;
            args.noise_count[c] = low_total + a_noise * total;
            args.total_count[c] = total;
            args.iterations[c] = it;
        }
    }
}
