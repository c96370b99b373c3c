// emStreamKernel — EM of ONE giant path cluster on the whole GPU, the matrix streamed from HBM every iteration.
//
// The sliced-ELL kernels (em_sell.cuh) keep a cluster in the shared memory of at most 16 SMs (one thread-block cluster);
// a path cluster far beyond that (hundreds of thousands of entries and more: deep metagenomic strain sets, a whole
// pantranscriptome left as one component) would crawl there. Here all SMs work on it: a cooperative launch of one
// persistent grid, the read classes and the paths each cut into one contiguous range per CTA of about equal entries,
// two grid-wide barriers per iteration:
//
//   E-step  rows of the CTA's range, cut once more into one contiguous range per warp. The entries of a warp's rows are ONE
//           contiguous stretch of the row-major arrays: the warp streams it through a ring of shared-memory stages with
//           asynchronous 16-byte copies (128 entries a stage, three stages always under way, whatever the rows look
//           like), multiplies a landed stage by the gathered abundances in place, and every lane sums its own read class
//           in entry order, 32 read classes at a time: w_i = count_i / (sum_j p_ij a_j + noise_i a_noise) into global
//           memory; the CTA's share of sum_i noise_i w_i into partial[cta]
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
constexpr uint32_t kStreamStep = 128;   // entries of one stage of a warp's ring
constexpr uint32_t kStreamStages = 4;

// One stage: 128 entries of the row-major arrays, aligned on a multiple of 128 entries (16-byte copies)
struct alignas(16) StreamStage {
    double val[kStreamStep];
    uint32_t col[kStreamStep];
};

constexpr size_t kStreamSmemBytes = sizeof(StreamStage) * kStreamStages * (kStreamThreads / 32);

// 16 bytes global -> shared without a register in between; bytes < 16: the rest is filled with zeros, 0: nothing is read
__device__ __forceinline__ void streamCopy16(void * dst, const void * src, uint32_t bytes) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(d), "l"(src), "r"(bytes) : "memory");
}

__device__ __forceinline__ void streamCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }

template <int N>
__device__ __forceinline__ void streamWait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

struct StreamArgs {

    EmArgs em;
    uint32_t cluster;
    double * partial;        // [grid] noise partial of every CTA
    uint32_t * flags;        // [grid] convergence vote of every CTA
    double * low_partial;    // [grid] mass below min_em_abundance (final pass)
};

__global__ void __launch_bounds__(kStreamThreads, 2) emStreamKernel(const StreamArgs sa) {

    __shared__ double red[32];
    extern __shared__ __align__(16) unsigned char stream_smem[];

    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();

    const EmArgs & args = sa.em;

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    StreamStage * ring = reinterpret_cast<StreamStage *>(stream_smem) + warp * kStreamStages;
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

    // ---- rows of this warp: the CTA's range cut the same way; their entries are the stretch [E0, E1) of the arrays
    uint32_t wrow_lo, wrow_hi;
    {
        constexpr uint32_t num_warps = kStreamThreads / 32;
        const uint32_t n = row_hi - row_lo;
        const uint64_t base = row_ptr[row_lo];
        const uint64_t total = (row_ptr[row_hi] - base) + n;

        auto cut = [&](uint32_t part) {
            if (part == 0) { return 0u; }
            if (part >= num_warps) { return n; }
            const uint64_t target = (total * part) / num_warps;
            uint32_t lo = 0, hi = n;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if ((row_ptr[row_lo + mid] - base) + mid >= target) { hi = mid; } else { lo = mid + 1; }
            }
            return lo;
        };
        wrow_lo = row_lo + cut(warp);
        wrow_hi = row_lo + cut(warp + 1);
    }

    const uint64_t E0 = row_ptr[wrow_lo];
    const uint64_t E1 = row_ptr[wrow_hi];
    const uint64_t cluster_end = e0 + Z;
    const uint64_t first_block = E0 & ~static_cast<uint64_t>(kStreamStep - 1);
    const uint32_t n_blocks = E1 > E0 ? static_cast<uint32_t>((E1 - first_block + kStreamStep - 1) / kStreamStep) : 0;

    // stage `slot` <- the 128 entries from `b` on (a multiple of 128: both arrays are read in aligned 16-byte pieces; what
    // lies before E0 or after E1 belongs to somebody else and is masked when the stage is used, what lies after the
    // cluster's last entry is not read)
    auto issue = [&](uint64_t b, uint32_t slot) {

        StreamStage & st = ring[slot];

        #pragma unroll
        for (uint32_t u = 0; u < 2; ++u) {
            const uint32_t q = u * 32 + lane;
            const uint64_t e = b + 2 * q;
            const uint32_t bytes = e + 2 <= cluster_end ? 16u : (e < cluster_end ? 8u : 0u);
            streamCopy16(st.val + 2 * q, args.val + (bytes ? e : b), bytes);
        }
        {
            const uint64_t e = b + 4 * lane;
            const uint32_t bytes = e + 4 <= cluster_end ? 16u : (e < cluster_end ? 4u * static_cast<uint32_t>(cluster_end - e) : 0u);
            streamCopy16(st.col + 4 * lane, args.col + (bytes ? e : b), bytes);
        }
    };

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
        // (the arrays are read once per iteration and go past the L1; a and w stay in the L2)
        #pragma unroll
        for (uint32_t p = 0; p + 1 < kStreamStages; ++p) {
            if (p < n_blocks) { issue(first_block + static_cast<uint64_t>(p) * kStreamStep, p); }
            streamCommit();
        }

        // the 32 read classes under way (lane = read class), and the 32 after them, loaded ahead
        uint32_t group = wrow_lo;
        uint64_t begin = 0, end = 0, next_begin = 0, next_end = 0;
        double nz = 0, ct = 0, next_nz = 0, next_ct = 0;

        if (group + lane < wrow_hi) {
            begin = row_ptr[group + lane];
            end = row_ptr[group + lane + 1];
            nz = noise[group + lane];
            ct = cnt[group + lane];
        }
        if (group + 32 + lane < wrow_hi) {
            next_begin = row_ptr[group + 32 + lane];
            next_end = row_ptr[group + 32 + lane + 1];
            next_nz = noise[group + 32 + lane];
            next_ct = cnt[group + 32 + lane];
        }

        double s = 0;

        // closes the group under way (all its entries are in) and moves on to the next one
        auto finish_group = [&]() {

            if (group + lane < wrow_hi) {
                const double wi = ct / (s + nz * a_noise);
                w[group + lane] = wi;
                noise_acc += nz * wi;
            }

            group += 32;
            begin = next_begin;
            end = next_end;
            nz = next_nz;
            ct = next_ct;
            s = 0;

            next_begin = next_end = 0;
            next_nz = next_ct = 0;
            if (group + 32 + lane < wrow_hi) {
                next_begin = row_ptr[group + 32 + lane];
                next_end = row_ptr[group + 32 + lane + 1];
                next_nz = noise[group + 32 + lane];
                next_ct = cnt[group + 32 + lane];
            }
        };

        for (uint32_t k = 0; k < n_blocks; ++k) {

            // (the stage refilled here was used up in the previous round)
            if (k + kStreamStages - 1 < n_blocks) { issue(first_block + static_cast<uint64_t>(k + kStreamStages - 1) * kStreamStep, (k + kStreamStages - 1) % kStreamStages); }
            streamCommit();
            streamWait<kStreamStages - 1>();
            __syncwarp();

            StreamStage & st = ring[k % kStreamStages];
            const uint64_t bs = first_block + static_cast<uint64_t>(k) * kStreamStep;
            const uint64_t be = min(bs + kStreamStep, E1);

            // products in place: all gathers of the stage under way together
            double v[kStreamStep / 32];
            uint32_t x[kStreamStep / 32];

            #pragma unroll
            for (uint32_t u = 0; u < kStreamStep / 32; ++u) {
                const uint64_t e = bs + u * 32 + lane;
                const bool mine = e >= E0 && e < E1;
                x[u] = mine ? st.col[u * 32 + lane] : 0u;
                v[u] = mine ? st.val[u * 32 + lane] : 0.0;
            }

            double g[kStreamStep / 32];

            #pragma unroll
            for (uint32_t u = 0; u < kStreamStep / 32; ++u) { g[u] = a[x[u]]; }

            #pragma unroll
            for (uint32_t u = 0; u < kStreamStep / 32; ++u) { st.val[u * 32 + lane] = v[u] * g[u]; }

            __syncwarp();

            // every lane adds the products of its own read class, in entry order; a group whose last entry is in this
            // stage is closed, the next one may begin in the same stage
            while (group < wrow_hi) {

                const uint64_t from = max(begin, bs);
                const uint64_t to = min(end, be);
                for (uint64_t e = from; e < to; ++e) { s += st.val[e - bs]; }

                const uint32_t last = min(31u, wrow_hi - 1 - group);
                const uint64_t group_end = __shfl_sync(0xffffffffu, end, last);
                if (group_end > be) { break; }

                finish_group();
            }

            __syncwarp();
        }

        // (read classes without entries after the warp's last entry, or a range without any entry)
        while (group < wrow_hi) { finish_group(); }

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
