// emSellKernel — the EM kernel proper (sm_100a).
//
// One thread-block cluster of S CTAs (S = 1, 2, 4, 8, 16) per path cluster; read classes (rows) are split into S
// slices of equal weight. In its prologue every CTA re-lays its slice out as two sliced-ELL structures, one per
// orientation, built so that the inner loops of both EM steps are the same branch-free, bank-conflict-free loop:
//
//   chunk = kSellSlots = 32 * kSellWays slots per step; a lane works on slots lane, lane + 32, ... (kSellWays
//   independent dependency chains per lane, which is what hides the shared-memory latency). A row (column) of the
//   chunk owns L = 1..32 adjacent slots (L chosen from its length so a chunk is ~8 steps deep); entry t of that row
//   sits at [chunk_off + (t / L) * kSellSlots + first_slot + (t % L)].
//   Rows (columns) are ordered by length (stable bucket sort) so rows sharing a chunk have near-equal length; padding
//   entries are (value 0, index 0) and need no predicate.
//
// E-step: warp per row chunk  : acc = sum val * a[col]; reduce over L lanes; w_row = count / (acc + noise * a_noise)
// M-step: warp per column chunk: acc = sum val * w[row]; reduce over L lanes; partial column sum g[col] (S > 1) or the
//         abundance update itself (S = 1).
// S > 1: reduce-scatter + all-gather of the N + 1 column sums over distributed shared memory, two cluster barriers per
//        iteration; S = 1: two block barriers per iteration (the second one carries the convergence vote).
//
// The slice lives in shared memory as far as it fits (vectors first, then column chunks, then row chunks); chunks
// that do not fit are streamed from a global scratch area with coalesced 128-byte warp loads ("hybrid residency").
// Everything is deterministic: fixed slices, stable sorts, fixed summation trees, no atomics.
#pragma once

#include <type_traits>

#include <cooperative_groups.h>

#include "common.cuh"
#include "em_kernels.cuh"

struct SellArgs {

    EmArgs em;
    const uint32_t * group_cluster;    // [n_groups] path cluster of each thread-block cluster
    const uint64_t * group_scratch;    // [n_groups] byte offset of the group's scratch area in `scratch`
    const uint64_t * group_cta_bytes;  // [n_groups] scratch bytes each CTA of the group owns
    unsigned char * scratch;
    uint32_t smem_bytes;               // dynamic shared memory per CTA of this launch
};

struct SellChunk {

    uint32_t off;     // first entry slot (multiple of 32)
    uint32_t first;   // first row (column) of the chunk, in sorted order
    uint16_t count;   // rows (columns) in the chunk, <= kSellSlots >> l2
    uint8_t l2;       // log2(lanes per row)
    uint8_t pad;
    uint32_t width;   // steps
};

constexpr uint32_t kSellBuckets = 64;

#ifndef RPVG_B200_SELL_WAYS
#define RPVG_B200_SELL_WAYS 1
#endif

#ifndef RPVG_B200_SELL_ROUNDS
#define RPVG_B200_SELL_ROUNDS 1   // depth divisor (see sellDepth): 1 = about two chunks per warp and phase
#endif

constexpr uint32_t kSellWays = RPVG_B200_SELL_WAYS;   // slots per lane and step
constexpr uint32_t kSellSlots = 32 * kSellWays;

// Length bucket: exact up to 32, then one bucket per power of two.
__host__ __device__ __forceinline__ uint32_t sellBucket(uint32_t len) {

    if (len <= 32) { return len; }
    uint32_t lg = 0;
    uint32_t v = len - 1;
    while (v >>= 1) { lg++; }   // floor(log2(len - 1)) >= 5
    return 32 + (lg - 4);
}

__host__ __device__ __forceinline__ uint32_t sellBucketUpper(uint32_t bucket) {

    return bucket <= 32 ? bucket : (1u << (bucket - 32 + 5));
}

// log2 of the lanes a row of (at most) this length gets: about `depth` entries per lane, 1..32 lanes.
__host__ __device__ __forceinline__ uint32_t sellLanesLog2(uint32_t len_upper, uint32_t depth) {

    const uint32_t want = (len_upper + depth - 1) / depth;
    uint32_t l2 = 0;
    while (l2 < 5 && (1u << l2) < want) { l2++; }
    return l2;
}

// Entry slots one orientation of a slice occupies (z entries in n rows): worst case and what is typical after sorting.
__host__ __device__ inline uint64_t sellSlotsWorst(uint64_t z, uint64_t n) { return 4 * z + 32 * n + 10 * kSellSlots * kSellWays + 3 * kSellSlots; }
__host__ __device__ inline uint64_t sellSlotsTypical(uint64_t z, uint64_t n) { return z + z / 4 + 2 * n + 4 * kSellSlots * kSellWays + 3 * kSellSlots; }

__host__ __device__ inline uint64_t sellRound16(uint64_t bytes) { return (bytes + 15) & ~uint64_t(15); }

// Shared-memory need of one CTA: m rows, z entries, N columns (worst: upper bound; otherwise the expected need).
__host__ __device__ inline uint64_t sellSmemBytes(uint64_t m, uint64_t z, uint64_t N, uint64_t index_bytes, bool multi, bool worst) {

    const uint64_t fixed = sellRound16(8 * (N + 1)) * (multi ? 2 : 1);
    const uint64_t vectors = 3 * sellRound16(8 * (m + 1));
    const uint64_t chunks = worst ? 16 * (m + N + 130) : 16 * ((m + N) / 8 + 140);
    const uint64_t col_id = sellRound16(index_bytes * (N + 1));
    const uint64_t slots = worst ? sellSlotsWorst(z, m) + sellSlotsWorst(z, N) : sellSlotsTypical(z, m) + sellSlotsTypical(z, N);
    const uint64_t masks = multi ? sellRound16(2 * (N + 2)) : 0;
    return fixed + vectors + chunks + col_id + masks + (8 + index_bytes) * slots + 256;
}

// Global scratch one CTA owns: prologue temporaries, everything that may spill out of shared memory and, when
// with_entries, the worst-case entry arrays.
__host__ __device__ inline uint64_t sellScratchBytes(uint64_t m, uint64_t z, uint64_t N, uint64_t index_bytes, bool with_entries) {

    const uint64_t temps = 2 * sellRound16(4 * (m + 1)) + 4 * sellRound16(4 * (N + 1));
    const uint64_t vectors = 3 * sellRound16(8 * (m + 1));
    const uint64_t chunks = 2 * 16 * (m + N + 130) + 2 * 4 * (m + N + 260);   // final and unassigned descriptors, assignment
    const uint64_t col_id = sellRound16(index_bytes * (N + 1));
    const uint64_t entries = with_entries ? (8 + index_bytes) * (sellSlotsWorst(z, m) + sellSlotsWorst(z, N)) + 128 : 64;
    const uint64_t masks = sellRound16(2 * (N + 2));
    return ((temps + vectors + chunks + col_id + masks + entries + 1023) / 1024) * 1024;
}

template <int THREADS>
struct SellSort {

    // shared state of the stable bucket sort (static shared memory)
    uint32_t hist[THREADS / 32][kSellBuckets];
    uint32_t bucket_total[kSellBuckets];
    uint32_t bucket_start[kSellBuckets];    // start of each bucket in sorted (descending length) order
    uint32_t chunk_base[kSellBuckets + 1];  // first chunk of each bucket
};

// Stable sort of n items by descending bucket. len(i) gives the length of item i. Writes pos[i] = sorted position and
// src[sorted position] = i. Warps own contiguous item ranges, so ties keep their original order.
template <int THREADS, typename LenFn>
__device__ void sellStableSort(SellSort<THREADS> & st, const uint32_t n, LenFn len, uint32_t * pos, uint32_t * src) {

    constexpr uint32_t W = THREADS / 32;
    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;

    for (uint32_t k = tid; k < W * kSellBuckets; k += THREADS) { (&st.hist[0][0])[k] = 0; }
    __syncthreads();

    const uint32_t per_warp = (((n + W - 1) / W) + 31) & ~31u;
    const uint32_t begin = min(n, warp * per_warp);
    const uint32_t end = min(n, begin + per_warp);

    for (uint32_t i0 = begin; i0 < end; i0 += 32) {

        const uint32_t i = i0 + lane;
        const uint32_t b = (i < end) ? sellBucket(len(i)) : 0xffffffffu;
        const uint32_t same = __match_any_sync(0xffffffffu, b);

        if (i < end && (same & ((1u << lane) - 1)) == 0) { st.hist[warp][b] += __popc(same); }
        __syncwarp();
    }

    __syncthreads();

    if (tid < kSellBuckets) {
        uint32_t total = 0;
        for (uint32_t w = 0; w < W; ++w) {
            const uint32_t h = st.hist[w][tid];
            st.hist[w][tid] = total;    // exclusive over warps
            total += h;
        }
        st.bucket_total[tid] = total;
    }

    __syncthreads();

    if (tid == 0) {
        uint32_t run = 0;
        for (int b = kSellBuckets - 1; b >= 0; --b) {
            st.bucket_start[b] = run;
            run += st.bucket_total[b];
        }
    }

    __syncthreads();

    for (uint32_t i0 = begin; i0 < end; i0 += 32) {

        const uint32_t i = i0 + lane;
        const uint32_t b = (i < end) ? sellBucket(len(i)) : 0xffffffffu;
        const uint32_t same = __match_any_sync(0xffffffffu, b);
        const uint32_t before = __popc(same & ((1u << lane) - 1));

        if (i < end) {
            const uint32_t p = st.bucket_start[b] + st.hist[warp][b] + before;
            pos[i] = p;
            src[p] = i;
        }
        __syncwarp();
        if (i < end && before == 0) { st.hist[warp][b] += __popc(same); }
        __syncwarp();
    }

    __syncthreads();
}

// Number of chunks sellMakeChunks will form for the items just sorted (same walk, nothing written).
// Entries per lane a side with z entries aims for: chunks cost ~100 instructions on top of ~6 per step, so they should
// be as deep as the warp count allows (about two rounds of chunks per warp), but at least 8 and at most 64 steps.
__host__ __device__ __forceinline__ uint32_t sellDepth(uint32_t z, uint32_t num_warps) {

    const uint32_t d = (2 * z) / (RPVG_B200_SELL_ROUNDS * kSellSlots * num_warps);
    return d < 8 ? 8 : (d > 64 ? 64 : d);
}

template <int THREADS>
__device__ uint32_t sellCountChunks(SellSort<THREADS> & st, const uint32_t n, const uint32_t depth) {

    if (threadIdx.x == 0) {

        uint32_t p = 0;
        uint32_t k = 0;
        int b = kSellBuckets - 1;

        while (p < n) {
            while (st.bucket_start[b] + st.bucket_total[b] <= p) { --b; }
            p += min(kSellSlots >> sellLanesLog2(sellBucketUpper(b), depth), n - p);
            ++k;
        }

        st.chunk_base[1] = k;
    }

    __syncthreads();
    return st.chunk_base[1];
}

// Chunk descriptors for items already sorted by descending bucket: a chunk takes the next kSellSlots >> l2 items, l2 from the
// bucket of its first (longest) item. Returns the number of chunks; total_slots receives the number of entry slots.
template <int THREADS, typename LenFn>
__device__ uint32_t sellMakeChunks(SellSort<THREADS> & st, const uint32_t n, const uint32_t depth, LenFn len, const uint32_t * src, SellChunk * chunks, uint32_t * scan_tmp, uint32_t * total_slots) {

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;

    if (tid == 0) {

        uint32_t p = 0;
        uint32_t k = 0;
        int b = kSellBuckets - 1;

        while (p < n) {

            while (st.bucket_start[b] + st.bucket_total[b] <= p) { --b; }

            const uint32_t l2 = sellLanesLog2(sellBucketUpper(b), depth);
            const uint32_t count = min(kSellSlots >> l2, n - p);

            SellChunk ch;
            ch.off = 0;
            ch.first = p;
            ch.count = static_cast<uint16_t>(count);
            ch.l2 = static_cast<uint8_t>(l2);
            ch.pad = 0;
            ch.width = 0;
            chunks[k] = ch;

            p += count;
            ++k;
        }

        st.chunk_base[0] = k;
    }

    __syncthreads();

    const uint32_t n_chunks = st.chunk_base[0];

    for (uint32_t k = tid; k < n_chunks; k += THREADS) {

        const uint32_t first = chunks[k].first;
        const uint32_t count = chunks[k].count;
        const uint32_t l2 = chunks[k].l2;

        uint32_t max_len = 0;
        for (uint32_t r = 0; r < count; ++r) { max_len = max(max_len, len(src[first + r])); }

        chunks[k].width = (max_len + (1u << l2) - 1) >> l2;
    }

    __syncthreads();

    // exclusive scan of width * kSellSlots over the chunks (block-wide, chunked per thread)
    const uint32_t per_thread = (n_chunks + THREADS - 1) / THREADS;
    const uint32_t k_begin = min(n_chunks, tid * per_thread);
    const uint32_t k_end = min(n_chunks, k_begin + per_thread);

    uint32_t local = 0;
    for (uint32_t k = k_begin; k < k_end; ++k) { local += chunks[k].width * kSellSlots; }

    uint32_t incl = local;
    #pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) { incl += up; }
    }
    if (lane == 31) { scan_tmp[warp] = incl; }
    __syncthreads();

    uint32_t warp_base = 0;
    for (uint32_t w = 0; w < warp; ++w) { warp_base += scan_tmp[w]; }

    uint32_t offset = warp_base + incl - local;
    for (uint32_t k = k_begin; k < k_end; ++k) {
        chunks[k].off = offset;
        offset += chunks[k].width * kSellSlots;
    }

    if (tid == THREADS - 1) { *total_slots = offset; }
    __syncthreads();

    return n_chunks;
}

// Deals the chunks to the warps: longest-processing-time-first on (width + a fixed per-chunk cost), so that every warp of
// the CTA has about the same number of steps in each EM phase. `in` holds the descriptors in the order they were made
// (widest first, roughly); `out` receives them grouped by warp, warp_begin[w] .. warp_begin[w + 1] being warp w's.
constexpr uint32_t kSellChunkCost = 5;   // steps a chunk costs on top of its width (descriptor, reduction, division)

template <int THREADS>
__device__ void sellAssignChunks(const SellChunk * in, const uint32_t n_chunks, SellChunk * out, uint32_t * warp_begin, uint32_t * dest) {

    constexpr uint32_t W = THREADS / 32;

    if (threadIdx.x == 0) {

        uint32_t load[W];
        uint32_t count[W];
        for (uint32_t w = 0; w < W; ++w) { load[w] = 0; count[w] = 0; }

        for (uint32_t k = 0; k < n_chunks; ++k) {

            uint32_t best = 0;
            for (uint32_t w = 1; w < W; ++w) {
                if (load[w] < load[best]) { best = w; }
            }

            load[best] += in[k].width + kSellChunkCost;
            count[best]++;
            dest[k] = best;
        }

        uint32_t run = 0;
        for (uint32_t w = 0; w < W; ++w) {
            warp_begin[w] = run;
            run += count[w];
            count[w] = warp_begin[w];
        }
        warp_begin[W] = run;

        for (uint32_t k = 0; k < n_chunks; ++k) { dest[k] = count[dest[k]]++; }
    }

    __syncthreads();

    for (uint32_t k = threadIdx.x; k < n_chunks; k += THREADS) { out[dest[k]] = in[k]; }

    __syncthreads();
}

// Bump allocator over the CTA's dynamic shared memory with overflow into its global scratch area.
struct SellArena {

    unsigned char * smem;
    uint64_t smem_left;
    unsigned char * gmem;
    uint64_t gmem_left;
    bool overflow;

    __device__ __forceinline__ void * take(uint64_t bytes, bool allow_smem, bool * in_smem) {

        bytes = (bytes + 15) & ~uint64_t(15);

        if (allow_smem && bytes <= smem_left) {
            void * p = smem;
            smem += bytes;
            smem_left -= bytes;
            if (in_smem) { *in_smem = true; }
            return p;
        }

        if (in_smem) { *in_smem = false; }
        if (bytes > gmem_left) {
            overflow = true;
            return gmem;   // never dereferenced: the kernel bails out on overflow
        }

        void * p = gmem;
        gmem += bytes;
        gmem_left -= bytes;
        return p;
    }
};

template <typename IndexT>
struct SellSide {

    const SellChunk * chunks;    // grouped by the warp that works on them (warp_begin)
    const uint32_t * warp_begin; // [num_warps + 1]
    uint32_t n_chunks;
    uint32_t n_resident;         // number of chunks whose entries are in shared memory: those with off < resident_slots
    uint32_t resident_slots;     // entry slots of the resident chunks
    const double * val_s;        // resident entries
    const IndexT * idx_s;
    const double * val_g;        // streamed entries (indexed by slot - resident_slots)
    const IndexT * idx_g;
};

template <typename IndexT>
struct SellLoop {

    double * a;
    double * g;
    double * w;
    const double * noise;
    const double * cnt;
    const IndexT * col_id;
    const uint16_t * owned_mask;
    SellSide<IndexT> rows;
    SellSide<IndexT> cols;
};

// 1 / x to within an ulp or two: hardware seed (20 bits) and two Newton steps. The E-step divides once per read class;
// an IEEE division would be a dependent chain four times as long on the critical path of every iteration.
__device__ __forceinline__ double sellReciprocal(double x) {

    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    return r;
}

// sum over the 1 << l2 adjacent lanes that share a row (column)
__device__ __forceinline__ double sellLaneSum(double acc, uint32_t l2) {

    if (l2 > 0) { acc += __shfl_xor_sync(0xffffffffu, acc, 1); }
    if (l2 > 1) { acc += __shfl_xor_sync(0xffffffffu, acc, 2); }
    if (l2 > 2) { acc += __shfl_xor_sync(0xffffffffu, acc, 4); }
    if (l2 > 3) { acc += __shfl_xor_sync(0xffffffffu, acc, 8); }
    if (l2 > 4) { acc += __shfl_xor_sync(0xffffffffu, acc, 16); }
    return acc;
}

// the same for the kSellWays partial sums of a lane at once (independent shuffle chains)
__device__ __forceinline__ void sellLaneSumWays(double (&acc)[kSellWays], uint32_t l2) {

    #pragma unroll
    for (uint32_t step = 0; step < 5; ++step) {
        if (l2 > step) {
            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) { acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], 1u << step); }
        }
    }
}

// sum_t val[t] * x[idx[t]] over the steps of a chunk for each of the lane's slots. (Measured on B200: the plain loop,
// one running sum per slot and four steps unrolled, beats grouped/predicated loads and split sums — the phase is bound by
// the shared-memory pipe, not by the dependent adds.)
template <typename IndexT>
__device__ __forceinline__ void sellChunkDot(const double * __restrict__ val, const IndexT * __restrict__ idx, const double * x, const uint32_t width, double (&acc)[kSellWays]) {

    #pragma unroll
    for (uint32_t u = 0; u < kSellWays; ++u) { acc[u] = 0; }

    #pragma unroll 4
    for (uint32_t t = 0; t < width; ++t) {
        #pragma unroll
        for (uint32_t u = 0; u < kSellWays; ++u) { acc[u] = fma(val[t * kSellSlots + u * 32], x[idx[t * kSellSlots + u * 32]], acc[u]); }
    }
}

#ifdef RPVG_B200_EM_TRACE
#define SELL_TRACE_DECL long long trace_t = clock64(); long long trace_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define SELL_TRACE(k) { const long long trace_now = clock64(); trace_acc[k] += trace_now - trace_t; trace_t = trace_now; }
#else
#define SELL_TRACE_DECL
#define SELL_TRACE(k)
#endif

// The EM iterations of one CTA (see the header comment of this file). Returns the number of iterations executed.
template <int THREADS, bool MULTI, bool ALL_SHARED, typename IndexT>
__device__ __forceinline__ uint32_t emSellIterate(const SellLoop<IndexT> & lp, const rpvg_b200_em_params & params, const uint32_t N, const uint32_t S, const uint32_t rank,
                                                  const uint32_t jc0, const uint32_t jc1, const double total, double * red, uint32_t * conv_flags) {

    cg::cluster_group cluster = cg::this_cluster();

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    constexpr uint32_t num_warps = THREADS / 32;

    double * a = lp.a;
    double * g = lp.g;
    double * w = lp.w;

    const double conv = params.max_rel_em_conv;
    const double min_abundance = params.min_em_abundance;
    const double inv_total = 1.0 / total;

    const uint32_t row_k0 = lp.rows.warp_begin[warp], row_k1 = lp.rows.warp_begin[warp + 1];
    const uint32_t col_k0 = lp.cols.warp_begin[warp], col_k1 = lp.cols.warp_begin[warp + 1];

    uint32_t conv_its = 0;
    uint32_t it = 0;

    SELL_TRACE_DECL

    while (it < params.max_em_its) {

        // ---- E-step over the slice: w_i = count_i / sum_j p_ij a_j
        const double a_noise = a[N];
        double noise_acc = 0;

        #pragma unroll 1
        for (uint32_t k = row_k0; k < row_k1; ++k) {

            const SellChunk ch = lp.rows.chunks[k];
            const bool lead = (lane & ((1u << ch.l2) - 1)) == 0;

            uint32_t p[kSellWays];
            bool head[kSellWays];
            double nz[kSellWays], ct[kSellWays];

            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) {
                const uint32_t r = (u * 32 + lane) >> ch.l2;
                head[u] = lead && r < ch.count;
                p[u] = ch.first + r;
                nz[u] = 0;
                ct[u] = 0;
                if (head[u]) {
                    nz[u] = lp.noise[p[u]];
                    ct[u] = lp.cnt[p[u]];
                }
            }

            const bool res = ALL_SHARED || ch.off < lp.rows.resident_slots;
            const double * val = (res ? lp.rows.val_s + ch.off : lp.rows.val_g + (ch.off - lp.rows.resident_slots)) + lane;
            const IndexT * idx = (res ? lp.rows.idx_s + ch.off : lp.rows.idx_g + (ch.off - lp.rows.resident_slots)) + lane;

            double acc[kSellWays];
            sellChunkDot<IndexT>(val, idx, a, ch.width, acc);
            sellLaneSumWays(acc, ch.l2);

            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) {
                if (head[u]) {
                    const double wv = ct[u] * sellReciprocal(fma(nz[u], a_noise, acc[u]));
                    w[p[u]] = wv;
                    noise_acc = fma(nz[u], wv, noise_acc);
                }
            }
        }

        SELL_TRACE(0)

        // noise column: per-warp partials now, combined in a fixed order after the barrier
        noise_acc = warpSum(noise_acc);
        if (lane == 0) { red[warp] = noise_acc; }
        if (MULTI && tid < S) { conv_flags[tid] = 1; }

        __syncthreads();

        SELL_TRACE(1)

        // ---- M-step over the slice: column sums (and, with one CTA, the abundance update itself)
        bool ok = true;

        #pragma unroll 1
        for (uint32_t k = col_k0; k < col_k1; ++k) {

            const SellChunk ch = lp.cols.chunks[k];
            const bool lead = (lane & ((1u << ch.l2) - 1)) == 0;

            uint32_t j[kSellWays];
            bool head[kSellWays];
            double a_old[kSellWays];

            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) {
                const uint32_t r = (u * 32 + lane) >> ch.l2;
                head[u] = lead && r < ch.count;
                j[u] = 0;
                a_old[u] = 0;
                if (head[u]) {
                    j[u] = lp.col_id[ch.first + r];
                    if (!MULTI) { a_old[u] = a[j[u]]; }
                }
            }

            const bool res = ALL_SHARED || ch.off < lp.cols.resident_slots;
            const double * val = (res ? lp.cols.val_s + ch.off : lp.cols.val_g + (ch.off - lp.cols.resident_slots)) + lane;
            const IndexT * idx = (res ? lp.cols.idx_s + ch.off : lp.cols.idx_g + (ch.off - lp.cols.resident_slots)) + lane;

            double acc[kSellWays];
            sellChunkDot<IndexT>(val, idx, w, ch.width, acc);
            sellLaneSumWays(acc, ch.l2);

            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) {
                if (head[u]) {
                    if (MULTI) {
                        g[j[u]] = acc[u];
                    } else {
                        const double a_new = (a_old[u] * acc[u]) * inv_total;
                        a[j[u]] = a_new;
                        if (a_new >= min_abundance && fabs(a_new - a_old[u]) > conv * a_new) { ok = false; }
                    }
                }
            }
        }

        if (tid == THREADS - 1) {

            double noise_g = 0;
            #pragma unroll
            for (uint32_t wi = 0; wi < num_warps; ++wi) { noise_g += red[wi]; }

            if (MULTI) {
                g[N] = noise_g;
            } else {
                const double a_new = (a_noise * noise_g) * inv_total;
                a[N] = a_new;
                if (a_new >= min_abundance && fabs(a_new - a_noise) > conv * a_new) { ok = false; }
            }
        }

        ++it;

        SELL_TRACE(2)

        if constexpr (!MULTI) {

            const int block_ok = __syncthreads_and(ok);
            SELL_TRACE(3)

            if (block_ok) {
                if (++conv_its == params.min_em_conv_its) { break; }
            } else {
                conv_its = 0;
            }

        } else {

            cluster.sync();

            SELL_TRACE(3)

            // reduce-scatter + all-gather over the group, restricted to the CTAs that hold the column: this CTA owns
            // columns [jc0, jc1); partial sums are added in ascending CTA order. The noise column (owned by the last
            // CTA, held by everyone) is reduced by one warp, a lane per CTA, so that its S remote loads overlap.
            const uint32_t jc1_paths = min(jc1, N);

            for (uint32_t j = jc0 + tid; j < jc1_paths; j += THREADS) {

                const uint32_t mask = lp.owned_mask[j - jc0];
                double sum = 0;

                if (__popc(mask) <= 2) {

                    for (uint32_t mm = mask; mm; mm &= mm - 1) { sum += *(cluster.map_shared_rank(g, __ffs(mm) - 1) + j); }

                } else {

                    double part[kMaxGroupSize];
                    #pragma unroll
                    for (uint32_t s = 0; s < kMaxGroupSize; ++s) { part[s] = ((mask >> s) & 1u) ? *(cluster.map_shared_rank(g, s) + j) : 0.0; }
                    #pragma unroll
                    for (uint32_t s = 0; s < kMaxGroupSize; ++s) { sum += part[s]; }
                }

                const double a_old = a[j];
                const double a_new = (a_old * sum) * inv_total;
                if (a_new >= min_abundance && fabs(a_new - a_old) > conv * a_new) { ok = false; }

                a[j] = a_new;
                for (uint32_t mm = mask & ~(1u << rank); mm; mm &= mm - 1) { *(cluster.map_shared_rank(a, __ffs(mm) - 1) + j) = a_new; }
            }

            if (jc1 == N + 1 && warp == num_warps - 1) {

                double part = (lane < S) ? *(cluster.map_shared_rank(g, lane) + N) : 0.0;
                part = warpSum(part);

                const double a_new = (a_noise * part) * inv_total;
                if (a_new >= min_abundance && fabs(a_new - a_noise) > conv * a_new) { ok = false; }

                if (lane < S) { *(cluster.map_shared_rank(a, lane) + N) = a_new; }
            }

            // convergence vote: flags were set to 1 before the block barrier above; a CTA in which any column failed
            // clears its flag in every CTA of the group
            SELL_TRACE(4)

            const int cta_failed = __syncthreads_or(!ok);
            if (cta_failed && tid < S) { *(cluster.map_shared_rank(conv_flags, tid) + rank) = 0; }

            cluster.sync();

            SELL_TRACE(5)

            bool all_ok = true;
            for (uint32_t s = 0; s < S; ++s) { all_ok = all_ok && (conv_flags[s] != 0); }

            if (all_ok) {
                if (++conv_its == params.min_em_conv_its) { break; }
            } else {
                conv_its = 0;
            }
        }
    }

#ifdef RPVG_B200_EM_TRACE
    if (lane == 0 && (warp == 0 || warp == num_warps - 1) && (rank == 0 || rank == S - 1) && it > 50) {
        printf("[trace] rank %u/%u warp %u its %u N %u chunks %u/%u cycles/it: E %lld | bar %lld | M %lld | sync1 %lld | reduce %lld | sync2 %lld\n", rank, S, warp, it, N,
               lp.rows.n_chunks, lp.cols.n_chunks, trace_acc[0] / it, trace_acc[1] / it, trace_acc[2] / it, trace_acc[3] / it, trace_acc[4] / it, trace_acc[5] / it);
    }
#endif

    return it;
}

#ifdef RPVG_B200_EM_TIMELINE
// tuning builds only: one record per CTA (tools/em_timeline.py)
__device__ unsigned long long rpvg_timeline_records[12 * 65536];
__device__ unsigned int rpvg_timeline_count;
#endif

#ifndef RPVG_B200_EM_MIN_CTAS_512
#define RPVG_B200_EM_MIN_CTAS_512 1
#endif

template <int THREADS, bool MULTI, typename IndexT>
__global__ void __launch_bounds__(THREADS, (THREADS == 512 ? RPVG_B200_EM_MIN_CTAS_512 : (THREADS == 128 ? 6 : (THREADS == 64 ? 12 : 1)))) emSellKernel(const SellArgs sa) {

    extern __shared__ __align__(16) unsigned char smem_raw[];

    __shared__ double red[32];
    __shared__ uint32_t conv_flags[kMaxGroupSize];
    __shared__ double final_part[kMaxGroupSize];
    __shared__ uint32_t scan_tmp[THREADS / 32];
    __shared__ uint32_t shared_u32[4];
    __shared__ uint32_t row_warp_begin[THREADS / 32 + 1];
    __shared__ uint32_t col_warp_begin[THREADS / 32 + 1];
    __shared__ SellSort<THREADS> sort_state;

    const EmArgs & args = sa.em;

    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t S = MULTI ? cluster.num_blocks() : 1;
    const uint32_t rank = MULTI ? cluster.block_rank() : 0;

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    constexpr uint32_t num_warps = THREADS / 32;

    if (*args.error) { return; }   // malformed arena: indices cannot be trusted (uniform over the whole grid)

#ifdef RPVG_B200_EM_TIMELINE
    unsigned long long timeline_t0, timeline_t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t0));
#endif

    const uint32_t group = blockIdx.x / S;
    const uint32_t c = sa.group_cluster[group];

    const uint64_t r0 = args.read_offset[c];
    const uint32_t M = static_cast<uint32_t>(args.read_offset[c + 1] - r0);
    const uint64_t p0 = args.path_offset[c];
    const uint32_t N = static_cast<uint32_t>(args.path_offset[c + 1] - p0);

    if (M == 0) {   // rpvg:src/path_abundance_estimator.cpp:20-22 — only the reset runs

        for (uint32_t j = rank * THREADS + tid; j < N; j += S * THREADS) { args.abundances[p0 + j] = 0; }

        if (rank == 0 && tid == 0) {
            args.noise_count[c] = 0;
            args.total_count[c] = 0;
            args.iterations[c] = 0;
        }
        return;
    }

    const uint64_t * row_ptr = args.row_ptr + r0;
    const uint64_t e_cluster = row_ptr[0];
    const uint64_t Z = row_ptr[M] - e_cluster;

    // ---- this CTA's slice of read classes: equal weight (entries + rows) per CTA
    uint32_t bounds[2];
    #pragma unroll
    for (int b = 0; b < 2; ++b) {

        const uint32_t r = rank + b;
        if (r == 0) { bounds[b] = 0; continue; }
        if (r >= S) { bounds[b] = M; continue; }

        const uint64_t target = ((Z + M) * r) / S;
        uint32_t lo = 0, hi = M;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if ((row_ptr[mid] - e_cluster) + mid >= target) { hi = mid; } else { lo = mid + 1; }
        }
        bounds[b] = lo;
    }

    const uint32_t i0 = bounds[0];
    const uint32_t m = bounds[1] - i0;
    const uint64_t e_slice = row_ptr[i0];

    // columns (noise included as column N) this CTA reduces and owns in the finalisation
    const uint32_t jc0 = static_cast<uint32_t>((static_cast<uint64_t>(N + 1) * rank) / S);
    const uint32_t jc1 = static_cast<uint32_t>((static_cast<uint64_t>(N + 1) * (rank + 1)) / S);

    // ---- memory: shared first, scratch for what does not fit
    SellArena arena;
    arena.smem = smem_raw;
    arena.smem_left = sa.smem_bytes;
    arena.gmem = sa.scratch + sa.group_scratch[group] + static_cast<uint64_t>(rank) * sa.group_cta_bytes[group];
    arena.gmem_left = sa.group_cta_bytes[group];
    arena.overflow = false;

    // a and g come first so that they sit at the same offsets in every CTA of the group (remote access); the host only
    // sends a cluster here when they fit
    double * a = static_cast<double *>(arena.take(8ull * (N + 1), true, nullptr));
    double * g = MULTI ? static_cast<double *>(arena.take(8ull * (N + 1), true, nullptr)) : nullptr;

    // prologue temporaries (global scratch only)
    uint32_t * row_pos = static_cast<uint32_t *>(arena.take(4ull * (m + 1), false, nullptr));
    uint32_t * row_src = static_cast<uint32_t *>(arena.take(4ull * (m + 1), false, nullptr));
    uint32_t * cstart = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * clen = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * col_pos = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * col_src = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));

    // per-row vectors in sorted order
    double * w = static_cast<double *>(arena.take(8ull * (m + 1), true, nullptr));
    double * noise = static_cast<double *>(arena.take(8ull * (m + 1), true, nullptr));
    double * cnt = static_cast<double *>(arena.take(8ull * (m + 1), true, nullptr));

    // ---- total read count: every CTA sums all rows in the same order, so all agree bit for bit
    double count_sum = 0;
    for (uint32_t i = tid; i < M; i += THREADS) { count_sum += args.cnt[r0 + i]; }
    const double total = blockSum(count_sum, red);

    // ---- rows: sort by length, chunk, fill
    auto row_len = [&](uint32_t i) { return static_cast<uint32_t>(row_ptr[i0 + i + 1] - row_ptr[i0 + i]); };
    sellStableSort<THREADS>(sort_state, m, row_len, row_pos, row_src);

    const uint32_t z_slice = static_cast<uint32_t>(row_ptr[i0 + m] - e_slice);
    const uint32_t depth = sellDepth(z_slice, num_warps);

    const uint32_t row_chunk_bound = sellCountChunks<THREADS>(sort_state, m, depth);

    SellChunk * row_chunks = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (row_chunk_bound + 1), true, nullptr));
    SellChunk * row_chunks_made = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (row_chunk_bound + 1), false, nullptr));
    uint32_t * row_chunk_dest = static_cast<uint32_t *>(arena.take(4ull * (row_chunk_bound + 1), false, nullptr));
    const uint32_t n_row_chunks = sellMakeChunks<THREADS>(sort_state, m, depth, row_len, row_src, row_chunks_made, scan_tmp, &shared_u32[0]);
    const uint32_t row_slots = shared_u32[0];

    for (uint32_t p = tid; p < m; p += THREADS) {
        const uint32_t i = row_src[p];
        noise[p] = args.noise[r0 + i0 + i];
        cnt[p] = args.cnt[r0 + i0 + i];
    }

    // ---- columns: range of this slice inside each column of the column-major arrays (rows ascend within a column)
    {
        const uint64_t * col_ptr = args.col_ptr + p0;
        const uint32_t i1 = i0 + m;

        for (uint32_t j = tid; j < N; j += THREADS) {

            const uint64_t k_begin = col_ptr[j];
            const uint64_t k_end = col_ptr[j + 1];

            uint64_t lo = k_begin, hi = k_end;
            while (lo < hi) {
                const uint64_t mid = (lo + hi) >> 1;
                if (__ldg(args.csc_row + mid) >= i0) { hi = mid; } else { lo = mid + 1; }
            }
            const uint64_t first = lo;

            hi = k_end;
            while (lo < hi) {
                const uint64_t mid = (lo + hi) >> 1;
                if (__ldg(args.csc_row + mid) >= i1) { hi = mid; } else { lo = mid + 1; }
            }

            cstart[j] = static_cast<uint32_t>(first);
            clen[j] = static_cast<uint32_t>(lo - first);
        }
    }

    __syncthreads();

    // ---- which CTAs of the group hold entries of each column this CTA owns: only those take part in the column's
    // reduction and need its abundance. With read classes ordered by their first path (em.cu) a slice touches a band of
    // columns, so most masks have one or two bits.
    uint16_t * owned_mask = nullptr;

    if (MULTI) {

        owned_mask = static_cast<uint16_t *>(arena.take(2ull * (jc1 - jc0 + 1), true, nullptr));

        unsigned char * present = reinterpret_cast<unsigned char *>(g);   // g is not in use yet
        for (uint32_t j = tid; j < N; j += THREADS) { present[j] = clen[j] > 0; }

        cluster.sync();

        for (uint32_t j = jc0 + tid; j < jc1; j += THREADS) {

            uint32_t mask = 0;
            if (j == N) {
                mask = (1u << S) - 1;   // the noise column concerns everyone
            } else {
                for (uint32_t s = 0; s < S; ++s) { mask |= static_cast<uint32_t>(*(reinterpret_cast<unsigned char *>(cluster.map_shared_rank(g, s)) + j)) << s; }
            }
            owned_mask[j - jc0] = static_cast<uint16_t>(mask);
        }

        cluster.sync();   // g may be overwritten from here on
    }

    auto col_len = [&](uint32_t j) { return clen[j]; };
    sellStableSort<THREADS>(sort_state, N, col_len, col_pos, col_src);

    // Only the columns with entries in this slice get chunks (empty ones sort last). A column that is empty in every
    // slice has no read class behind it: its abundance is 0 from the first update on (a_j * 0 / total), which is what it
    // is set to — with one CTA right away (nothing ever reads it), with several by the reduction (mask 0, sum 0).
    const uint32_t n_cols_present = N - sort_state.bucket_total[0];

    const uint32_t col_chunk_bound = sellCountChunks<THREADS>(sort_state, n_cols_present, depth);

    SellChunk * col_chunks = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (col_chunk_bound + 1), true, nullptr));
    SellChunk * col_chunks_made = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (col_chunk_bound + 1), false, nullptr));
    uint32_t * col_chunk_dest = static_cast<uint32_t *>(arena.take(4ull * (col_chunk_bound + 1), false, nullptr));
    const uint32_t n_col_chunks = sellMakeChunks<THREADS>(sort_state, n_cols_present, depth, col_len, col_src, col_chunks_made, scan_tmp, &shared_u32[1]);
    const uint32_t col_slots = shared_u32[1];

    // column ids in sorted order (kept for the whole kernel)
    IndexT * col_id = static_cast<IndexT *>(arena.take(sizeof(IndexT) * (n_cols_present + 1), true, nullptr));
    for (uint32_t p = tid; p < n_cols_present; p += THREADS) { col_id[p] = static_cast<IndexT>(col_src[p]); }

    // ---- entry arrays: column side first (it also carries the update), then row side; whole chunks only
    auto place = [&](const SellChunk * chunks, uint32_t n_chunks, uint32_t slots, SellSide<IndexT> * side) {

        // largest prefix of chunks whose entries fit in what is left of shared memory
        const uint64_t per_slot = 8 + sizeof(IndexT);
        const uint64_t budget_slots = arena.smem_left > 64 ? (arena.smem_left - 64) / per_slot : 0;

        uint32_t n_res = n_chunks;
        uint32_t res_slots = slots;

        if (slots > budget_slots) {
            uint32_t lo = 0, hi = n_chunks;   // first chunk whose end exceeds the budget
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (static_cast<uint64_t>(chunks[mid].off) + chunks[mid].width * kSellSlots > budget_slots) { hi = mid; } else { lo = mid + 1; }
            }
            n_res = lo;
            res_slots = (n_res < n_chunks) ? chunks[n_res].off : slots;
        }

        side->chunks = nullptr;   // set once the chunks are dealt to the warps
        side->n_chunks = n_chunks;
        side->n_resident = n_res;
        side->resident_slots = res_slots;
        side->val_s = static_cast<double *>(arena.take(8ull * res_slots, true, nullptr));
        side->idx_s = static_cast<IndexT *>(arena.take(sizeof(IndexT) * static_cast<uint64_t>(res_slots), true, nullptr));
        side->val_g = static_cast<double *>(arena.take(8ull * (slots - res_slots), false, nullptr));
        side->idx_g = static_cast<IndexT *>(arena.take(sizeof(IndexT) * static_cast<uint64_t>(slots - res_slots), false, nullptr));
    };

    SellSide<IndexT> cols, rows;
    place(col_chunks_made, n_col_chunks, col_slots, &cols);
    place(row_chunks_made, n_row_chunks, row_slots, &rows);

    // capacity check, uniform over the group: every CTA publishes its flag before anyone proceeds
    if (MULTI) {
        if (tid == 0) { conv_flags[0] = 0; }
        cluster.sync();
        if (arena.overflow && tid == 0) {
            for (uint32_t s = 0; s < S; ++s) { *(cluster.map_shared_rank(conv_flags, s)) = 1; }
            printf("[rpvg_b200] capacity: cluster %u rank %u/%u M %u N %u m %u row_slots %u col_slots %u chunks %u/%u cta_bytes %llu smem %u\n", c, rank, S, M, N, m,
                   row_slots, col_slots, n_row_chunks, n_col_chunks, static_cast<unsigned long long>(sa.group_cta_bytes[group]), sa.smem_bytes);
        }
        cluster.sync();
        if (conv_flags[0]) {
            if (rank == 0 && tid == 0) { atomicOr(const_cast<uint32_t *>(args.error), RPVG_DEV_ERR_CAPACITY); }
            return;
        }
    } else if (arena.overflow) {
        if (tid == 0) {
            atomicOr(const_cast<uint32_t *>(args.error), RPVG_DEV_ERR_CAPACITY);
            printf("[rpvg_b200] capacity: cluster %u M %u N %u m %u row_slots %u col_slots %u chunks %u/%u cta_bytes %llu smem %u\n", c, M, N, m, row_slots, col_slots,
                   n_row_chunks, n_col_chunks, static_cast<unsigned long long>(sa.group_cta_bytes[group]), sa.smem_bytes);
        }
        return;
    }

    // ---- deal the chunks to the warps (after the capacity check: the scratch arrays are known to be in bounds)
    sellAssignChunks<THREADS>(row_chunks_made, n_row_chunks, row_chunks, row_warp_begin, row_chunk_dest);
    sellAssignChunks<THREADS>(col_chunks_made, n_col_chunks, col_chunks, col_warp_begin, col_chunk_dest);

    rows.chunks = row_chunks;
    rows.warp_begin = row_warp_begin;
    cols.chunks = col_chunks;
    cols.warp_begin = col_warp_begin;

    // ---- fill: one warp per chunk
    auto fill = [&](const SellSide<IndexT> & side, auto entry_range, auto entry_value, auto entry_index, const uint32_t * src) {

        for (uint32_t k = warp; k < side.n_chunks; k += num_warps) {

            const SellChunk ch = side.chunks[k];
            const uint32_t lanes = 1u << ch.l2;
            const uint32_t sub = lane & (lanes - 1);

            const bool res = ch.off < side.resident_slots;
            double * val = const_cast<double *>(res ? side.val_s + ch.off : side.val_g + (ch.off - side.resident_slots));
            IndexT * idx = const_cast<IndexT *>(res ? side.idx_s + ch.off : side.idx_g + (ch.off - side.resident_slots));

            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) {

                const uint32_t slot = u * 32 + lane;
                const uint32_t r = slot >> ch.l2;

                uint64_t base = 0;
                uint32_t len = 0;
                if (r < ch.count) { entry_range(src[ch.first + r], &base, &len); }

                for (uint32_t t = 0; t < ch.width; ++t) {

                    const uint32_t e = t * lanes + sub;
                    double v = 0;
                    IndexT x = 0;
                    if (e < len) {
                        v = entry_value(base + e);
                        x = static_cast<IndexT>(entry_index(base + e));
                    }
                    val[t * kSellSlots + slot] = v;
                    idx[t * kSellSlots + slot] = x;
                }
            }
        }
    };

    fill(rows,
         [&](uint32_t i, uint64_t * base, uint32_t * len) { *base = row_ptr[i0 + i]; *len = static_cast<uint32_t>(row_ptr[i0 + i + 1] - *base); },
         [&](uint64_t e) { return args.val[e]; },
         [&](uint64_t e) { return args.col[e]; },
         row_src);

    fill(cols,
         [&](uint32_t j, uint64_t * base, uint32_t * len) { *base = cstart[j]; *len = clen[j]; },
         [&](uint64_t k) { return args.csc_val[k]; },
         [&](uint64_t k) { return row_pos[args.csc_row[k] - i0]; },
         col_src);

    // rpvg:src/path_abundance_estimator.cpp:54 — single-precision quotient widened to double
    const double a_init = static_cast<double>(1 / static_cast<float>(N + 1));
    for (uint32_t j = tid; j <= N; j += THREADS) { a[j] = (MULTI || j == N || clen[j] > 0) ? a_init : 0.0; }

    if (MULTI) { cluster.sync(); } else { __syncthreads(); }

    // ---- EM iterations. When everything this CTA touches sits in shared memory the loop runs on pointers the compiler
    // can prove to be shared (plain LDS/STS); otherwise on generic pointers (hybrid residency).
    SellLoop<IndexT> loop;
    loop.a = a;
    loop.g = g;
    loop.w = w;
    loop.noise = noise;
    loop.cnt = cnt;
    loop.col_id = col_id;
    loop.owned_mask = owned_mask;
    loop.rows = rows;
    loop.cols = cols;

    const unsigned char * smem_end = smem_raw + sa.smem_bytes;
    auto in_shared = [&](const void * ptr) {
        const unsigned char * q = static_cast<const unsigned char *>(ptr);
        return ptr == nullptr || (q >= smem_raw && q < smem_end);
    };

    const bool all_shared = in_shared(a) && in_shared(g) && in_shared(w) && in_shared(noise) && in_shared(cnt) && in_shared(col_id) && in_shared(owned_mask) &&
                            in_shared(rows.chunks) && in_shared(cols.chunks) && rows.n_resident == rows.n_chunks && cols.n_resident == cols.n_chunks;

    const double total_count = total;
    uint32_t it;

#ifdef RPVG_B200_EM_TRACE
    if (tid == 0 && (rank == 0 || rank == S - 1)) {
        uint32_t widest_row = 0, widest_col = 0;
        for (uint32_t k = 0; k < n_row_chunks; ++k) { widest_row = max(widest_row, row_chunks[k].width); }
        for (uint32_t k = 0; k < n_col_chunks; ++k) { widest_col = max(widest_col, col_chunks[k].width); }
        printf("[layout] rank %u/%u m %u z %u cols present %u | rows: chunks %u slots %u widest %u resident %u | cols: chunks %u slots %u widest %u resident %u | depth %u\n", rank, S, m, z_slice,
               n_cols_present, n_row_chunks, row_slots, widest_row, rows.n_resident, n_col_chunks, col_slots, widest_col, cols.n_resident, depth);
    }
#endif

#ifdef RPVG_B200_EM_TIMELINE
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t1));
#endif

    if (all_shared) {

        auto rebase = [&](auto * ptr) {
            using T = std::remove_pointer_t<decltype(ptr)>;
            if (ptr == nullptr) { return static_cast<T *>(nullptr); }
            const uint32_t off = static_cast<uint32_t>(reinterpret_cast<const unsigned char *>(ptr) - smem_raw);
            return reinterpret_cast<T *>(smem_raw + off);
        };

        SellLoop<IndexT> fast = loop;
        fast.a = rebase(loop.a);
        fast.g = rebase(loop.g);
        fast.w = rebase(loop.w);
        fast.noise = rebase(loop.noise);
        fast.cnt = rebase(loop.cnt);
        fast.col_id = rebase(loop.col_id);
        fast.owned_mask = rebase(loop.owned_mask);
        fast.rows.chunks = rebase(loop.rows.chunks);
        fast.rows.val_s = rebase(loop.rows.val_s);
        fast.rows.idx_s = rebase(loop.rows.idx_s);
        fast.cols.chunks = rebase(loop.cols.chunks);
        fast.cols.val_s = rebase(loop.cols.val_s);
        fast.cols.idx_s = rebase(loop.cols.idx_s);

        it = emSellIterate<THREADS, MULTI, true, IndexT>(fast, args.params, N, S, rank, jc0, jc1, total_count, red, conv_flags);

    } else {

        it = emSellIterate<THREADS, MULTI, false, IndexT>(loop, args.params, N, S, rank, jc0, jc1, total_count, red, conv_flags);
    }

    const double min_abundance = args.params.min_em_abundance;


    // rpvg:src/path_abundance_estimator.cpp:100-113; every CTA finalises the columns it owns
    double noise_part = 0;
    for (uint32_t j = jc0 + tid; j < min(jc1, N); j += THREADS) {
        const double aj = a[j];
        if (aj < min_abundance) {
            noise_part += aj * total;
            args.abundances[p0 + j] = 0;
        } else {
            args.abundances[p0 + j] = aj * total;
        }
    }

    const double low_sum = blockSum(noise_part, red);

    if (MULTI) {
        if (tid == 0) { *(cluster.map_shared_rank(final_part, 0) + rank) = low_sum; }
        cluster.sync();
    }

    if (rank == 0 && tid == 0) {

        double low_total = low_sum;
        if (MULTI) {
            low_total = 0;
            for (uint32_t s = 0; s < S; ++s) { low_total += final_part[s]; }
        }

        // the noise column is owned by the last CTA of the group; every CTA holds its value
        args.noise_count[c] = low_total + a[N] * total;
        args.total_count[c] = total;
        args.iterations[c] = it;
    }

#ifdef RPVG_B200_EM_TIMELINE
    if (tid == 0) {
        unsigned long long timeline_t2;
        uint32_t smid;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t2));
        asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        const unsigned int slot = atomicAdd(&rpvg_timeline_count, 1u);
        if (slot < 65536) {
            unsigned long long * rec = rpvg_timeline_records + 12ull * slot;
            rec[0] = THREADS; rec[1] = S; rec[2] = rank; rec[3] = c; rec[4] = smid; rec[5] = M; rec[6] = N;
            rec[7] = timeline_t0; rec[8] = timeline_t1; rec[9] = timeline_t2; rec[10] = it; rec[11] = all_shared ? 1u : 0u;
        }
    }
#endif

    // no CTA may exit while a peer can still touch its shared memory
    if (MULTI) { cluster.sync(); }
}
