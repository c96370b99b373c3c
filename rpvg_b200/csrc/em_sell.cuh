// emSellBuildKernel + emSellRunKernel — the EM kernels proper (sm_100a).
//
// One thread-block cluster of S CTAs (S = 1, 2, 4, 8, 16) per path cluster; read classes (rows) are split into S
// slices of equal weight. The build kernel re-lays every CTA's slice out as two sliced-ELL structures, one per
// orientation, built so that the inner loops of both EM steps are the same branch-free loop; the run kernel (same
// grid, same clusters, half the registers) iterates on them:
//
//   chunk = kSellSlots = 32 * kSellWays slots per step; a lane works on slots lane, lane + 32, ... (kSellWays = 1:
//   wider chunks were measured slower, the phases are bound by the shared-memory pipe). A row (column) of the
//   chunk owns L = 1..32 adjacent slots (L chosen from its length so a chunk is ~8 steps deep); entry t of that row
//   sits at [chunk_off + (t / L) * kSellSlots + first_slot + (t % L)].
//   Rows (columns) are ordered by length (stable bucket sort) so rows sharing a chunk have near-equal length; padding
//   entries are (value 0, index 0) and need no predicate.
//
// E-step: warp per row chunk  : acc = sum val * a[col]; reduce over L lanes; w_row = count / (acc + noise * a_noise)
// M-step: warp per column chunk: acc = sum val * w[row]; reduce over L lanes; partial column sum g[col] (S > 1) or the
//         abundance update itself (S = 1).
// S > 1: every CTA works on local ids of the columns its slice touches. After the M-step and ONE cluster barrier each
//        CTA pulls, for each of its columns, the partial sums of the other CTAs that hold the column (distributed shared
//        memory; the read classes are ordered by first path, so most columns live in one or two slices) and updates its
//        own copy of the abundance — all holders add in the same order and get the same bits. The convergence vote
//        travels one iteration behind (previous abundances are kept), so it costs no barrier either.
//        S = 1: two block barriers per iteration (the second one carries the convergence vote).
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
    uint32_t run_warps;                // warps of the CTAs that will iterate (the layout is made for them)
};

// Bytes at the end of a CTA's scratch area that hand the slice from the build kernel to the run kernel: a header and
// the image of the build kernel's dynamic shared memory.
constexpr uint64_t kSellHeaderBytes = 1024;
__host__ __device__ inline uint64_t sellHandoverBytes(uint64_t smem_bytes) { return kSellHeaderBytes + ((smem_bytes + 255) & ~uint64_t(255)); }

struct alignas(16) SellChunk {   // one 16-byte load

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

#ifndef RPVG_B200_SELL_SPREAD_MIN_THREADS
#define RPVG_B200_SELL_SPREAD_MIN_THREADS 256   // CTAs from this size on reorder their entries against bank conflicts
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

// Shared-memory need of one CTA: m rows, z entries, N columns of which n_cols have entries in the slice (worst: upper
// bound; otherwise the expected need). One CTA keeps the abundances of all N + 1 columns; a CTA of a group keeps
// vectors over its own columns only (two generations of column sums, abundances, previous abundances) and the tables of
// the exchange with its peers.
__host__ __device__ inline uint64_t sellSmemBytes(uint64_t m, uint64_t z, uint64_t N, uint64_t n_cols, uint64_t index_bytes, bool multi, bool worst) {

    const uint64_t fixed = multi ? 4 * sellRound16(8 * (n_cols + 2)) : sellRound16(8 * (N + 1));
    const uint64_t vectors = 3 * sellRound16(8 * (m + 1));
    const uint64_t chunks = worst ? 16 * (m + n_cols + 130) : 16 * ((m + n_cols) / 8 + 140);
    const uint64_t col_id = multi ? 0 : sellRound16(index_bytes * (N + 1));
    const uint64_t slots = worst ? sellSlotsWorst(z, m) + sellSlotsWorst(z, n_cols) : sellSlotsTypical(z, m) + sellSlotsTypical(z, n_cols);
    const uint64_t tables = multi ? (6 * (n_cols + 2) + (worst ? 2 * 15 * n_cols : n_cols) + 64) : 0;
    return fixed + vectors + chunks + col_id + tables + (8 + index_bytes) * slots + 256;
}

// Global scratch one CTA owns: prologue temporaries, everything that may spill out of shared memory and, when
// with_entries, the worst-case entry arrays.
__host__ __device__ inline uint64_t sellScratchBytes(uint64_t m, uint64_t z, uint64_t N, uint64_t index_bytes, bool with_entries) {

    const uint64_t temps = 2 * sellRound16(4 * (m + 1)) + 5 * sellRound16(4 * (N + 1));
    const uint64_t vectors = 3 * sellRound16(8 * (m + 1)) + 2 * sellRound16(8 * (N + 2));
    const uint64_t chunks = 2 * 16 * (m + N + 130) + 2 * 4 * (m + N + 260);   // final and unassigned descriptors, assignment
    const uint64_t col_id = sellRound16(index_bytes * (N + 1));
    const uint64_t entries = with_entries ? (8 + index_bytes) * (sellSlotsWorst(z, m) + sellSlotsWorst(z, N)) + 128 : 64;
    const uint64_t tables = 6 * (N + 2) + 2 * 15 * (N + 1) + 64;
    return ((temps + vectors + chunks + col_id + tables + entries + 1023) / 1024) * 1024;
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

// Exclusive prefix sum over n values held in (global or shared) memory, in place; returns the total. Block-wide; tmp has
// one slot per warp.
template <int THREADS>
__device__ uint32_t sellExclusiveScan(uint32_t * values, const uint32_t n, uint32_t * tmp) {

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;

    const uint32_t per_thread = (n + THREADS - 1) / THREADS;
    const uint32_t begin = min(n, tid * per_thread);
    const uint32_t end = min(n, begin + per_thread);

    uint32_t local = 0;
    for (uint32_t k = begin; k < end; ++k) { local += values[k]; }

    uint32_t incl = local;
    #pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) { incl += up; }
    }

    __syncthreads();   // tmp may still be read from an earlier call
    if (lane == 31) { tmp[warp] = incl; }
    __syncthreads();

    uint32_t warp_base = 0;
    uint32_t total = 0;
    for (uint32_t w = 0; w < THREADS / 32; ++w) {
        if (w < warp) { warp_base += tmp[w]; }
        total += tmp[w];
    }

    uint32_t run = warp_base + incl - local;
    for (uint32_t k = begin; k < end; ++k) {
        const uint32_t v = values[k];
        values[k] = run;
        run += v;
    }

    __syncthreads();
    return total;
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
__device__ void sellAssignChunks(const SellChunk * in, const uint32_t n_chunks, SellChunk * out, uint32_t * warp_begin, uint32_t * dest, const uint32_t W) {

    if (threadIdx.x == 0) {

        uint32_t load[32];
        uint32_t count[32];
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

// Exchange state of a CTA that is part of a group (static shared memory, same offsets in every CTA of the group).
struct SellPeers {

    uint32_t present[kMaxGroupSize];   // columns (with entries) of every CTA's slice: its noise column has that local id
    double noise_in[2][kMaxGroupSize]; // noise partial of every CTA, pushed before the barrier; one set per generation
    uint32_t flag_in[2][kMaxGroupSize];// "all my columns converged in the update before", likewise
    uint32_t vote[32];                 // per warp, from the update just done
    uint32_t stop;
};

template <typename IndexT>
struct SellLoop {

    double * a;                  // abundances: all N + 1 columns (one CTA) or this slice's columns + noise (group)
    double * a_prev;             // group: abundances before the latest update
    double * g;                  // group: two generations of partial column sums, g + gen * g_stride
    uint32_t g_stride;
    uint32_t n_local;            // group: columns of this slice; the noise column has local id n_local
    double * w;
    const double * noise;
    const double * cnt;
    const IndexT * col_id;       // one CTA: column of each sorted position
    const uint16_t * holders;    // group: bit s set = CTA s holds the column
    const uint32_t * pull_begin; // group: first entry of the column in peer_id
    const IndexT * peer_id;      // group: local id of the column in each other holder, ascending CTA rank
    SellPeers * peers;
    SellSide<IndexT> rows;
    SellSide<IndexT> cols;
};

// What the build kernel leaves for the run kernel at the end of the CTA's scratch area (sellHandoverBytes).
template <typename IndexT>
struct SellHeader {

    uint32_t state;            // 0: nothing to iterate (cluster without read classes, or capacity error); 1: slice ready
    uint32_t all_shared;       // everything the loop touches is in the shared-memory image
    uint32_t image_bytes;      // used part of the build kernel's dynamic shared memory, copied behind this header
    uint32_t n_cols_present;
    uint64_t smem_base;        // generic address of the build kernel's dynamic shared memory: pointers into
                               // [smem_base, smem_base + image_bytes) are relative to it
    double total;
    const uint32_t * col_src;  // column of each sorted position (global scratch)
    uint32_t row_warp_begin[33];
    uint32_t col_warp_begin[33];
    uint32_t present[kMaxGroupSize];
    SellLoop<IndexT> loop;
};

static_assert(sizeof(SellHeader<uint32_t>) <= kSellHeaderBytes, "SellHeader outgrew its slot");

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

// sum over the 1 << l2 adjacent lanes that share a row (column), for the kSellWays partial sums of a lane at once
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
// the shared-memory pipe, not by the dependent adds; eight steps in flight for chunks that stream from L2 lose as well.)
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

// E-step over the slice: w_i = count_i / (sum_j p_ij a_j + noise_i a_noise); returns this warp's share of
// sum_i noise_i w_i (summed over the warp).
template <bool ALL_SHARED, typename IndexT>
__device__ __forceinline__ double sellExpectation(const SellLoop<IndexT> & lp, const double * a, const double a_noise, const uint32_t k0, const uint32_t k1, const uint32_t lane) {

    double noise_acc = 0;

    #pragma unroll 1
    for (uint32_t k = k0; k < k1; ++k) {

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
                lp.w[p[u]] = wv;
                noise_acc = fma(nz[u], wv, noise_acc);
            }
        }
    }

    return warpSum(noise_acc);
}

// The EM iterations of a CTA that has the whole path cluster (see the header comment of this file). Returns the number of
// iterations executed.
template <int THREADS, bool ALL_SHARED, typename IndexT>
__device__ __forceinline__ uint32_t emSellIterateSingle(const SellLoop<IndexT> & lp, const rpvg_b200_em_params & params, const uint32_t N, const double total, double * red) {

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    constexpr uint32_t num_warps = THREADS / 32;

    double * a = lp.a;
    const double * w = lp.w;

    const double conv = params.max_rel_em_conv;
    const double min_abundance = params.min_em_abundance;
    const double inv_total = 1.0 / total;

    const uint32_t row_k0 = lp.rows.warp_begin[warp], row_k1 = lp.rows.warp_begin[warp + 1];
    const uint32_t col_k0 = lp.cols.warp_begin[warp], col_k1 = lp.cols.warp_begin[warp + 1];

    uint32_t conv_its = 0;
    uint32_t it = 0;

    SELL_TRACE_DECL

    while (it < params.max_em_its) {

        const double a_noise = a[N];
        const double noise_acc = sellExpectation<ALL_SHARED, IndexT>(lp, a, a_noise, row_k0, row_k1, lane);

        SELL_TRACE(0)

        // noise column: per-warp partials now, combined in a fixed order after the barrier
        if (lane == 0) { red[warp] = noise_acc; }

        __syncthreads();

        SELL_TRACE(1)

        // ---- M-step: column sums and the abundance update itself
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
                    a_old[u] = a[j[u]];
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
                    const double a_new = (a_old[u] * acc[u]) * inv_total;
                    a[j[u]] = a_new;
                    if (a_new >= min_abundance && fabs(a_new - a_old[u]) > conv * a_new) { ok = false; }
                }
            }
        }

        if (tid == THREADS - 1) {

            double noise_g = 0;
            #pragma unroll
            for (uint32_t wi = 0; wi < num_warps; ++wi) { noise_g += red[wi]; }

            const double a_new = (a_noise * noise_g) * inv_total;
            a[N] = a_new;
            if (a_new >= min_abundance && fabs(a_new - a_noise) > conv * a_new) { ok = false; }
        }

        ++it;

        SELL_TRACE(2)

        const int block_ok = __syncthreads_and(ok);

        SELL_TRACE(3)

        if (block_ok) {
            if (++conv_its == params.min_em_conv_its) { break; }
        } else {
            conv_its = 0;
        }
    }

#ifdef RPVG_B200_EM_TRACE
    if (lane == 0 && (warp == 0 || warp == num_warps - 1) && it > 50) {
        printf("[trace] one CTA warp %u its %u N %u chunks %u/%u cycles/it: E %lld | bar %lld | M %lld | vote %lld\n", warp, it, N, lp.rows.n_chunks, lp.cols.n_chunks,
               trace_acc[0] / it, trace_acc[1] / it, trace_acc[2] / it, trace_acc[3] / it);
    }
#endif

    return it;
}

// The EM iterations of one CTA of a group (thread-block cluster of S CTAs). Returns the number of iterations the
// estimate went through; *final_a receives the vector that holds it (a, or a_prev when the stop came one iteration late).
template <int THREADS, bool ALL_SHARED, typename IndexT>
__device__ __forceinline__ uint32_t emSellIterateGroup(const SellLoop<IndexT> & lp, const rpvg_b200_em_params & params, const uint32_t S, const uint32_t rank, const double total,
                                                       double * red, const double ** final_a) {

    cg::cluster_group cluster = cg::this_cluster();

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31;
    const uint32_t warp = tid >> 5;
    constexpr uint32_t num_warps = THREADS / 32;

    double * a = lp.a;
    double * a_prev = lp.a_prev;
    const double * w = lp.w;
    SellPeers * peers = lp.peers;

    const uint32_t n_local = lp.n_local;

    const double conv = params.max_rel_em_conv;
    const double min_abundance = params.min_em_abundance;
    const double inv_total = 1.0 / total;

    const uint32_t row_k0 = lp.rows.warp_begin[warp], row_k1 = lp.rows.warp_begin[warp + 1];
    const uint32_t col_k0 = lp.cols.warp_begin[warp], col_k1 = lp.cols.warp_begin[warp + 1];

    uint32_t conv_its = 0;   // kept by the last warp (uniform over the group: everyone sees the same flags)
    uint32_t it = 0;
    uint32_t gen = 0;
    bool stopped_late = false;

    SELL_TRACE_DECL

    while (it < params.max_em_its) {

        const double a_noise = a[n_local];
        const double noise_acc = sellExpectation<ALL_SHARED, IndexT>(lp, a, a_noise, row_k0, row_k1, lane);

        SELL_TRACE(0)

        if (lane == 0) { red[warp] = noise_acc; }

        __syncthreads();

        SELL_TRACE(1)

        // ---- M-step: this slice's share of the column sums, generation `gen`
        double * g = lp.g + gen * lp.g_stride;

        // The noise partial and this CTA's vote on the update before this one go to every CTA of the group right away:
        // plain remote stores, complete by the time the barrier lets anyone read them.
        if (warp == num_warps - 1) {

            const double noise_g = warpSum(lane < num_warps ? red[lane] : 0.0);
            const bool all = __all_sync(0xffffffffu, lane < num_warps ? peers->vote[lane] != 0 : true);

            if (lane < S) {
                *(cluster.map_shared_rank(&peers->noise_in[gen][rank], lane)) = noise_g;
                *(cluster.map_shared_rank(&peers->flag_in[gen][rank], lane)) = all ? 1u : 0u;
            }
        }

        #pragma unroll 1
        for (uint32_t k = col_k0; k < col_k1; ++k) {

            const SellChunk ch = lp.cols.chunks[k];
            const bool lead = (lane & ((1u << ch.l2) - 1)) == 0;

            const bool res = ALL_SHARED || ch.off < lp.cols.resident_slots;
            const double * val = (res ? lp.cols.val_s + ch.off : lp.cols.val_g + (ch.off - lp.cols.resident_slots)) + lane;
            const IndexT * idx = (res ? lp.cols.idx_s + ch.off : lp.cols.idx_g + (ch.off - lp.cols.resident_slots)) + lane;

            double acc[kSellWays];
            sellChunkDot<IndexT>(val, idx, w, ch.width, acc);
            sellLaneSumWays(acc, ch.l2);

            #pragma unroll
            for (uint32_t u = 0; u < kSellWays; ++u) {
                const uint32_t r = (u * 32 + lane) >> ch.l2;
                if (lead && r < ch.count) { g[ch.first + r] = acc[u]; }   // local id = sorted position
            }
        }

        ++it;

        SELL_TRACE(2)

        cluster.sync();   // the only group-wide barrier of the iteration

        SELL_TRACE(3)

        // ---- update: every holder of a column adds the partial sums of all its holders, in ascending CTA order
        bool ok = true;

        for (uint32_t l = tid; l < n_local; l += THREADS) {

            const uint32_t mask = lp.holders[l];
            double sum;

            if (mask == (1u << rank)) {

                sum = g[l];

            } else {

                uint32_t k = lp.pull_begin[l];
                sum = 0;
                for (uint32_t mm = mask; mm; mm &= mm - 1) {
                    const uint32_t s = __ffs(mm) - 1;
                    sum += (s == rank) ? g[l] : *(cluster.map_shared_rank(g, s) + lp.peer_id[k++]);
                }
            }

            const double a_old = a[l];
            const double a_new = (a_old * sum) * inv_total;
            if (a_new >= min_abundance && fabs(a_new - a_old) > conv * a_new) { ok = false; }

            a_prev[l] = a_old;
            a[l] = a_new;
        }

        if (warp == num_warps - 1) {

            // noise column: a lane per CTA, the same summation tree everywhere
            const double part = warpSum(lane < S ? peers->noise_in[gen][lane] : 0.0);
            const uint32_t peer_flag = (it > 1 && lane < S) ? peers->flag_in[gen][lane] : 1u;

            const double a_new = (a_noise * part) * inv_total;
            if (a_new >= min_abundance && fabs(a_new - a_noise) > conv * a_new) { ok = false; }

            if (lane == 0) {
                a_prev[n_local] = a_noise;
                a[n_local] = a_new;
            }

            // the group's verdict on the previous update (iteration it - 1)
            if (it > 1) {

                const bool all_ok = __all_sync(0xffffffffu, peer_flag != 0);
                uint32_t stop = 0;

                if (all_ok) {
                    if (++conv_its == params.min_em_conv_its) { stop = 1; }
                } else {
                    conv_its = 0;
                }
                if (lane == 0) { peers->stop = stop; }

            } else if (lane == 0) {
                peers->stop = 0;
            }
        }

        const bool warp_ok = __all_sync(0xffffffffu, ok);
        if (lane == 0) { peers->vote[warp] = warp_ok ? 1u : 0u; }

        SELL_TRACE(4)

        __syncthreads();   // abundances, votes and the stop flag are in place

        SELL_TRACE(5)

        if (peers->stop) {
            // iteration it - 1 was the one that completed the run of converged iterations: its result is a_prev
            stopped_late = true;
            --it;
            break;
        }

        gen ^= 1;
    }

#ifdef RPVG_B200_EM_TRACE
    if (lane == 0 && (warp == 0 || warp == num_warps - 1) && (rank == 0 || rank == S - 1) && it > 50) {
        printf("[trace] rank %u/%u warp %u its %u local cols %u chunks %u/%u cycles/it: E %lld | bar %lld | M %lld | cluster %lld | update %lld | bar2 %lld\n", rank, S, warp, it, n_local,
               lp.rows.n_chunks, lp.cols.n_chunks, trace_acc[0] / it, trace_acc[1] / it, trace_acc[2] / it, trace_acc[3] / it, trace_acc[4] / it, trace_acc[5] / it);
    }
#endif

    *final_a = stopped_late ? a_prev : a;
    return it;
}

#ifdef RPVG_B200_EM_TIMELINE
// tuning builds only: one record per CTA (tools/em_timeline.py)
__device__ unsigned long long rpvg_timeline_records[12 * 65536];
__device__ unsigned int rpvg_timeline_count;
#endif

// Build kernel: re-lays the slice out (see the header comment of this file) in its own shared memory and scratch area and
// hands the result to emSellRunKernel. Separate kernels because the layout code needs twice the registers of the EM loop:
// the loop alone fits 64 registers, which doubles the warps an SM holds while it iterates.
template <int THREADS, bool MULTI, typename IndexT>
__global__ void __launch_bounds__(THREADS, (THREADS == 128 ? 6 : (THREADS == 64 ? 12 : 1))) emSellBuildKernel(const SellArgs sa) {

    extern __shared__ __align__(16) unsigned char smem_raw[];

    __shared__ double red[32];
    __shared__ double final_part[kMaxGroupSize];
    __shared__ SellPeers peers;
    __shared__ uint32_t capacity_flag;
    __shared__ uint32_t scan_tmp[THREADS / 32];
    __shared__ uint32_t shared_u32[4];
    __shared__ uint32_t row_warp_begin[33];
    __shared__ uint32_t col_warp_begin[33];
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

    const uint32_t group = blockIdx.x / S;
    const uint32_t c = sa.group_cluster[group];

    unsigned char * cta_scratch = sa.scratch + sa.group_scratch[group] + static_cast<uint64_t>(rank) * sa.group_cta_bytes[group];
    const uint64_t cta_scratch_bytes = sa.group_cta_bytes[group] - sellHandoverBytes(sa.smem_bytes);
    SellHeader<IndexT> * header = reinterpret_cast<SellHeader<IndexT> *>(cta_scratch + cta_scratch_bytes);

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
        if (tid == 0) { header->state = 0; }
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

    // ---- memory: shared first, scratch for what does not fit
    SellArena arena;
    arena.smem = smem_raw;
    arena.smem_left = sa.smem_bytes;
    arena.gmem = cta_scratch;
    arena.gmem_left = cta_scratch_bytes;
    arena.overflow = false;

    const unsigned char * scratch_base = arena.gmem;

    // prologue temporaries (global scratch only); those sized by N come first, so that they sit at the same offsets in
    // the scratch areas of all CTAs of the group (col_pos is read by the peers)
    uint32_t * cstart = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * clen = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * col_pos = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * col_src = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * col_holders = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * col_pulls = static_cast<uint32_t *>(arena.take(4ull * (N + 1), false, nullptr));
    uint32_t * row_pos = static_cast<uint32_t *>(arena.take(4ull * (m + 1), false, nullptr));
    uint32_t * row_src = static_cast<uint32_t *>(arena.take(4ull * (m + 1), false, nullptr));

    // ---- total read count: every CTA sums all rows in the same order, so all agree bit for bit
    double count_sum = 0;
    for (uint32_t i = tid; i < M; i += THREADS) { count_sum += args.cnt[r0 + i]; }
    const double total = blockSum(count_sum, red);

    // ---- rows: sort by length, chunk (descriptors in scratch until they are dealt to the warps)
    auto row_len = [&](uint32_t i) { return static_cast<uint32_t>(row_ptr[i0 + i + 1] - row_ptr[i0 + i]); };
    sellStableSort<THREADS>(sort_state, m, row_len, row_pos, row_src);

    const uint32_t z_slice = static_cast<uint32_t>(row_ptr[i0 + m] - e_slice);
    const uint32_t depth = sellDepth(z_slice, sa.run_warps);

    const uint32_t row_chunk_bound = sellCountChunks<THREADS>(sort_state, m, depth);

    SellChunk * row_chunks_made = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (row_chunk_bound + 1), false, nullptr));
    uint32_t * row_chunk_dest = static_cast<uint32_t *>(arena.take(4ull * (row_chunk_bound + 1), false, nullptr));
    const uint32_t n_row_chunks = sellMakeChunks<THREADS>(sort_state, m, depth, row_len, row_src, row_chunks_made, scan_tmp, &shared_u32[0]);
    const uint32_t row_slots = shared_u32[0];

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

    // Columns sorted by their length in this slice; only those with entries get chunks (empty ones sort last). A column
    // that is empty in every slice has no read class behind it: its abundance is 0 from the first update on
    // (a_j * 0 / total).
    auto col_len = [&](uint32_t j) { return clen[j]; };
    sellStableSort<THREADS>(sort_state, N, col_len, col_pos, col_src);

    const uint32_t n_cols_present = N - sort_state.bucket_total[0];

#ifndef RPVG_B200_SELL_COL_DEPTH_DIV
#define RPVG_B200_SELL_COL_DEPTH_DIV 1
#endif
    const uint32_t col_depth = max(8u, depth / RPVG_B200_SELL_COL_DEPTH_DIV);

    const uint32_t col_chunk_bound = sellCountChunks<THREADS>(sort_state, n_cols_present, col_depth);

    SellChunk * col_chunks_made = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (col_chunk_bound + 1), false, nullptr));
    uint32_t * col_chunk_dest = static_cast<uint32_t *>(arena.take(4ull * (col_chunk_bound + 1), false, nullptr));
    const uint32_t n_col_chunks = sellMakeChunks<THREADS>(sort_state, n_cols_present, col_depth, col_len, col_src, col_chunks_made, scan_tmp, &shared_u32[1]);
    const uint32_t col_slots = shared_u32[1];

    // ---- group: local column ids (= sorted position), and for each of this slice's columns the other CTAs that hold it.
    // Nothing has been placed in shared memory yet: its start serves as exchange area.
    double * a = nullptr;
    double * a_prev = nullptr;
    double * g = nullptr;
    uint32_t g_stride = 0;
    uint32_t total_pulls = 0;

    if (MULTI) {

        // columns nobody holds: abundance 0, written by the CTA whose share of the path range they fall in
        const uint32_t jc0 = static_cast<uint32_t>((static_cast<uint64_t>(N) * rank) / S);
        const uint32_t jc1 = static_cast<uint32_t>((static_cast<uint64_t>(N) * (rank + 1)) / S);
        for (uint32_t j = jc0 + tid; j < jc1; j += THREADS) { args.abundances[p0 + j] = 0; }

        unsigned char * present = smem_raw;   // N flags, needed until the masks are made (N <= smem: see sellSmemBytes)
        for (uint32_t j = tid; j < N; j += THREADS) { present[j] = clen[j] > 0; }

        if (tid < S) { *(cluster.map_shared_rank(&peers.present[rank], tid)) = n_cols_present; }
        if (tid < 32) { peers.vote[tid] = 0; }
        if (tid == 0) { peers.stop = 0; }

        __threadfence();   // col_pos travels through L2
        cluster.sync();

        uint32_t most = 0;
        for (uint32_t s = 0; s < S; ++s) { most = max(most, peers.present[s]); }
        g_stride = (most + 2) & ~1u;   // columns + noise, even

        for (uint32_t l = tid; l < n_cols_present; l += THREADS) {

            const uint32_t j = col_src[l];
            uint32_t mask = 0;
            for (uint32_t s = 0; s < S; ++s) { mask |= static_cast<uint32_t>(*(cluster.map_shared_rank(present, s) + j)) << s; }

            col_holders[l] = mask;
            col_pulls[l] = __popc(mask) - 1;
        }

        cluster.sync();   // the flags may be overwritten from here on

        total_pulls = sellExclusiveScan<THREADS>(col_pulls, n_cols_present, scan_tmp);

        // the column sums are read by the peers: shared memory, same offset everywhere, or the group cannot run
        bool g_in_smem = false;
        g = static_cast<double *>(arena.take(16ull * g_stride, true, &g_in_smem));
        if (!g_in_smem) { arena.overflow = true; }

        a = static_cast<double *>(arena.take(8ull * g_stride, true, nullptr));
        a_prev = static_cast<double *>(arena.take(8ull * g_stride, true, nullptr));

    } else {

        a = static_cast<double *>(arena.take(8ull * (N + 1), true, nullptr));
    }

    uint16_t * holders = MULTI ? static_cast<uint16_t *>(arena.take(2ull * (n_cols_present + 1), true, nullptr)) : nullptr;
    uint32_t * pull_begin = MULTI ? static_cast<uint32_t *>(arena.take(4ull * (n_cols_present + 2), true, nullptr)) : nullptr;
    IndexT * peer_id = MULTI ? static_cast<IndexT *>(arena.take(sizeof(IndexT) * (static_cast<uint64_t>(total_pulls) + 1), true, nullptr)) : nullptr;

    // per-row vectors in sorted order
    double * w = static_cast<double *>(arena.take(8ull * (m + 1), true, nullptr));
    double * noise = static_cast<double *>(arena.take(8ull * (m + 1), true, nullptr));
    double * cnt = static_cast<double *>(arena.take(8ull * (m + 1), true, nullptr));

    // chunk descriptors as the loops will read them, and (one CTA) the column of each sorted position
    SellChunk * row_chunks = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (n_row_chunks + 1), true, nullptr));
    SellChunk * col_chunks = static_cast<SellChunk *>(arena.take(sizeof(SellChunk) * (n_col_chunks + 1), true, nullptr));
    IndexT * col_id = MULTI ? nullptr : static_cast<IndexT *>(arena.take(sizeof(IndexT) * (n_cols_present + 1), true, nullptr));

    // ---- entry arrays: column side first, then row side; whole chunks only
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
        side->warp_begin = nullptr;
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
        if (tid == 0) { capacity_flag = 0; }
        cluster.sync();
        if (arena.overflow && tid == 0) {
            for (uint32_t s = 0; s < S; ++s) { *(cluster.map_shared_rank(&capacity_flag, s)) = 1; }
            printf("[rpvg_b200] capacity: cluster %u rank %u/%u M %u N %u m %u columns %u (stride %u) row_slots %u col_slots %u chunks %u/%u pulls %u cta_bytes %llu smem %u\n", c, rank, S, M,
                   N, m, n_cols_present, g_stride, row_slots, col_slots, n_row_chunks, n_col_chunks, total_pulls, static_cast<unsigned long long>(sa.group_cta_bytes[group]), sa.smem_bytes);
        }
        cluster.sync();
        if (capacity_flag) {
            if (rank == 0 && tid == 0) { atomicOr(const_cast<uint32_t *>(args.error), RPVG_DEV_ERR_CAPACITY); }
            if (tid == 0) { header->state = 0; }
            return;
        }
    } else if (arena.overflow) {
        if (tid == 0) {
            atomicOr(const_cast<uint32_t *>(args.error), RPVG_DEV_ERR_CAPACITY);
            printf("[rpvg_b200] capacity: cluster %u M %u N %u m %u row_slots %u col_slots %u chunks %u/%u cta_bytes %llu smem %u\n", c, M, N, m, row_slots, col_slots,
                   n_row_chunks, n_col_chunks, static_cast<unsigned long long>(sa.group_cta_bytes[group]), sa.smem_bytes);
            header->state = 0;
        }
        return;
    }

    // ---- deal the chunks to the warps (after the capacity check: the scratch arrays are known to be in bounds)
    sellAssignChunks<THREADS>(row_chunks_made, n_row_chunks, row_chunks, row_warp_begin, row_chunk_dest, sa.run_warps);
    sellAssignChunks<THREADS>(col_chunks_made, n_col_chunks, col_chunks, col_warp_begin, col_chunk_dest, sa.run_warps);

    rows.chunks = row_chunks;
    rows.warp_begin = row_warp_begin;
    cols.chunks = col_chunks;
    cols.warp_begin = col_warp_begin;

    // ---- per-row vectors, in sorted order
    for (uint32_t p = tid; p < m; p += THREADS) {
        const uint32_t i = row_src[p];
        noise[p] = args.noise[r0 + i0 + i];
        cnt[p] = args.cnt[r0 + i0 + i];
    }

    // ---- group: exchange tables. The local id a column has in a peer is its sorted position there (the peer's col_pos).
    if (MULTI) {

        const uint64_t col_pos_offset = reinterpret_cast<const unsigned char *>(col_pos) - scratch_base;
        const unsigned char * group_base = sa.scratch + sa.group_scratch[group];
        const uint64_t cta_bytes = sa.group_cta_bytes[group];

        for (uint32_t l = tid; l < n_cols_present; l += THREADS) {

            const uint32_t mask = col_holders[l];
            const uint32_t j = col_src[l];
            uint32_t k = col_pulls[l];

            holders[l] = static_cast<uint16_t>(mask);
            pull_begin[l] = k;

            for (uint32_t mm = mask & ~(1u << rank); mm; mm &= mm - 1) {
                const uint32_t s = __ffs(mm) - 1;
                const uint32_t * peer_col_pos = reinterpret_cast<const uint32_t *>(group_base + s * cta_bytes + col_pos_offset);
                peer_id[k++] = static_cast<IndexT>(__ldcg(peer_col_pos + j));
            }
        }

        if (tid == 0) { pull_begin[n_cols_present] = total_pulls; }

    } else {

        for (uint32_t p = tid; p < n_cols_present; p += THREADS) { col_id[p] = static_cast<IndexT>(col_src[p]); }
    }

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

    // row side: the index is the column's local id in a group, the path's index in the cluster otherwise
    fill(rows,
         [&](uint32_t i, uint64_t * base, uint32_t * len) { *base = row_ptr[i0 + i]; *len = static_cast<uint32_t>(row_ptr[i0 + i + 1] - *base); },
         [&](uint64_t e) { return args.val[e]; },
         [&](uint64_t e) { return MULTI ? col_pos[args.col[e]] : args.col[e]; },
         row_src);

    fill(cols,
         [&](uint32_t j, uint64_t * base, uint32_t * len) { *base = cstart[j]; *len = clen[j]; },
         [&](uint64_t k) { return args.csc_val[k]; },
         [&](uint64_t k) { return row_pos[args.csc_row[k] - i0]; },
         col_src);

    // ---- spread the gathers over the banks. The EM phases are bound by the shared-memory pipe and nearly half of its
    // wavefronts are the fp64 gathers a[col] / w[row]: the 16 lanes of a half-warp read 16 eight-byte words, and words
    // whose index agrees mod 16 collide (4.2 wavefronts per gather in the E-step, 5.5 in the M-step as laid out; 2 is the
    // floor). The order of a lane's entries is free, so each lane permutes its own: step by step the lanes of a half-warp
    // pick, one after the other, an entry of theirs whose bank is still free (tools/bank_sim2.py: ~3.6 both ways).
    // Only for the larger CTAs: it costs a few thousand instructions per chunk.
    if (THREADS >= RPVG_B200_SELL_SPREAD_MIN_THREADS) {

        __syncwarp();

        auto spread = [&](const SellSide<IndexT> & side) {

            for (uint32_t k = warp; k < side.n_chunks; k += num_warps) {

                const SellChunk ch = side.chunks[k];
                if (ch.width < 2) { continue; }

                const bool res = ch.off < side.resident_slots;
                double * val = const_cast<double *>(res ? side.val_s + ch.off : side.val_g + (ch.off - side.resident_slots)) + lane;
                IndexT * idx = const_cast<IndexT *>(res ? side.idx_s + ch.off : side.idx_g + (ch.off - side.resident_slots)) + lane;

                const uint32_t half_base = lane & 16;

                for (uint32_t t = 0; t + 1 < ch.width; ++t) {

                    uint32_t taken = 0;
                    uint32_t pick = t;

                    for (uint32_t i = 0; i < 16; ++i) {

                        uint32_t bank_bit = 0;

                        if ((lane & 15) == i) {

                            uint32_t s_found = t;
                            for (uint32_t q = t; q < ch.width; ++q) {
                                if (((taken >> (static_cast<uint32_t>(idx[q * kSellSlots]) & 15)) & 1u) == 0) {
                                    s_found = q;
                                    break;
                                }
                            }
                            pick = s_found;
                            bank_bit = 1u << (static_cast<uint32_t>(idx[pick * kSellSlots]) & 15);
                        }

                        taken |= __shfl_sync(0xffffffffu, bank_bit, half_base + i);
                    }

                    if (pick != t) {
                        const double v = val[t * kSellSlots];
                        const IndexT x = idx[t * kSellSlots];
                        val[t * kSellSlots] = val[pick * kSellSlots];
                        idx[t * kSellSlots] = idx[pick * kSellSlots];
                        val[pick * kSellSlots] = v;
                        idx[pick * kSellSlots] = x;
                    }
                }
            }
        };

        spread(rows);
        spread(cols);
    }

    // rpvg:src/path_abundance_estimator.cpp:54 — single-precision quotient widened to double
    const double a_init = static_cast<double>(1 / static_cast<float>(N + 1));

    if (MULTI) {
        for (uint32_t l = tid; l <= n_cols_present; l += THREADS) {
            a[l] = a_init;
            a_prev[l] = a_init;
        }
        cluster.sync();
    } else {
        for (uint32_t j = tid; j <= N; j += THREADS) { a[j] = (j == N || clen[j] > 0) ? a_init : 0.0; }
        __syncthreads();
    }

    // ---- hand the slice over: loop arguments, and the image of this CTA's shared memory
    SellLoop<IndexT> loop;
    loop.a = a;
    loop.a_prev = a_prev;
    loop.g = g;
    loop.g_stride = g_stride;
    loop.n_local = n_cols_present;
    loop.w = w;
    loop.noise = noise;
    loop.cnt = cnt;
    loop.col_id = col_id;
    loop.holders = holders;
    loop.pull_begin = pull_begin;
    loop.peer_id = peer_id;
    loop.peers = nullptr;
    loop.rows = rows;
    loop.cols = cols;
    loop.rows.warp_begin = nullptr;
    loop.cols.warp_begin = nullptr;

    const unsigned char * smem_end = smem_raw + sa.smem_bytes;
    auto in_shared = [&](const void * ptr) {
        const unsigned char * q = static_cast<const unsigned char *>(ptr);
        return ptr == nullptr || (q >= smem_raw && q < smem_end);
    };

    const bool all_shared = in_shared(a) && in_shared(a_prev) && in_shared(g) && in_shared(w) && in_shared(noise) && in_shared(cnt) && in_shared(col_id) && in_shared(holders) &&
                            in_shared(pull_begin) && in_shared(peer_id) && in_shared(rows.chunks) && in_shared(cols.chunks) && rows.n_resident == rows.n_chunks &&
                            cols.n_resident == cols.n_chunks;

    const uint32_t image_bytes = static_cast<uint32_t>(arena.smem - smem_raw);

#ifdef RPVG_B200_EM_TRACE
    if (tid == 0 && (rank == 0 || rank == S - 1)) {
        uint32_t widest_row = 0, widest_col = 0;
        for (uint32_t k = 0; k < n_row_chunks; ++k) { widest_row = max(widest_row, row_chunks[k].width); }
        for (uint32_t k = 0; k < n_col_chunks; ++k) { widest_col = max(widest_col, col_chunks[k].width); }
        printf("[layout] rank %u/%u m %u z %u cols present %u pulls %u | rows: chunks %u slots %u widest %u resident %u | cols: chunks %u slots %u widest %u resident %u | depth %u all shared %d image %u\n",
               rank, S, m, z_slice, n_cols_present, total_pulls, n_row_chunks, row_slots, widest_row, rows.n_resident, n_col_chunks, col_slots, widest_col, cols.n_resident, depth, all_shared ? 1 : 0,
               image_bytes);
    }
#endif

    {
        const uint4 * src = reinterpret_cast<const uint4 *>(smem_raw);
        uint4 * dst = reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(header) + kSellHeaderBytes);
        for (uint32_t i = tid; i < (image_bytes + 15) / 16; i += THREADS) { dst[i] = src[i]; }
    }

    if (tid < 33) {
        header->row_warp_begin[tid] = tid <= sa.run_warps ? row_warp_begin[tid] : 0;
        header->col_warp_begin[tid] = tid <= sa.run_warps ? col_warp_begin[tid] : 0;
    }
    if (MULTI && tid < kMaxGroupSize) { header->present[tid] = tid < S ? peers.present[tid] : 0; }

    if (tid == 0) {
        header->all_shared = all_shared ? 1 : 0;
        header->image_bytes = image_bytes;
        header->n_cols_present = n_cols_present;
        header->smem_base = reinterpret_cast<uint64_t>(smem_raw);
        header->total = total;
        header->col_src = col_src;
        header->loop = loop;
        header->state = 1;
    }
}

template <int THREADS>
constexpr int sellRunMinBlocks() { return THREADS >= 1024 ? 1 : 1024 / THREADS; }   // 64 registers per thread

// Run kernel: takes the slice the build kernel left (same grid, same thread-block clusters), loads it into shared memory
// and iterates. Its register budget is half the build kernel's, so an SM holds 32 warps of it.
template <int THREADS, bool MULTI, typename IndexT>
__global__ void __launch_bounds__(THREADS, sellRunMinBlocks<THREADS>()) emSellRunKernel(const SellArgs sa) {

    extern __shared__ __align__(16) unsigned char smem_raw[];

    __shared__ double red[32];
    __shared__ double final_part[kMaxGroupSize];
    __shared__ SellPeers peers;
    __shared__ uint32_t row_warp_begin[33];
    __shared__ uint32_t col_warp_begin[33];

    const EmArgs & args = sa.em;

    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t S = MULTI ? cluster.num_blocks() : 1;
    const uint32_t rank = MULTI ? cluster.block_rank() : 0;

    const uint32_t tid = threadIdx.x;

    if (*args.error) { return; }   // malformed arena or a slice without room (uniform over the whole grid)

#ifdef RPVG_B200_EM_TIMELINE
    unsigned long long timeline_t0, timeline_t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t0));
#endif

    const uint32_t group = blockIdx.x / S;
    const uint32_t c = sa.group_cluster[group];

    const uint64_t p0 = args.path_offset[c];
    const uint32_t N = static_cast<uint32_t>(args.path_offset[c + 1] - p0);

    unsigned char * cta_scratch = sa.scratch + sa.group_scratch[group] + static_cast<uint64_t>(rank) * sa.group_cta_bytes[group];
    const SellHeader<IndexT> * header = reinterpret_cast<const SellHeader<IndexT> *>(cta_scratch + sa.group_cta_bytes[group] - sellHandoverBytes(sa.smem_bytes));

    if (header->state == 0) { return; }   // uniform over the group

    // ---- the slice: image into shared memory, loop arguments with their pointers moved along
    {
        const uint4 * src = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(header) + kSellHeaderBytes);
        uint4 * dst = reinterpret_cast<uint4 *>(smem_raw);
        const uint32_t n16 = (header->image_bytes + 15) / 16;
        for (uint32_t i = tid; i < n16; i += THREADS) { dst[i] = src[i]; }
    }

    if (tid < 33) {
        row_warp_begin[tid] = header->row_warp_begin[tid];
        col_warp_begin[tid] = header->col_warp_begin[tid];
    }
    if (tid < kMaxGroupSize) { peers.present[tid] = header->present[tid]; }
    if (tid < 32) { peers.vote[tid] = 0; }
    if (tid == 0) { peers.stop = 0; }

    SellLoop<IndexT> loop = header->loop;

    const uint64_t old_base = header->smem_base;
    const uint64_t image_bytes = header->image_bytes;

    auto moved = [&](auto * ptr) {
        using T = std::remove_pointer_t<decltype(ptr)>;
        const uint64_t address = reinterpret_cast<uint64_t>(ptr);
        if (ptr != nullptr && address >= old_base && address < old_base + image_bytes) { return reinterpret_cast<T *>(smem_raw + (address - old_base)); }
        return ptr;
    };

    loop.a = moved(loop.a);
    loop.a_prev = moved(loop.a_prev);
    loop.g = moved(loop.g);
    loop.w = moved(loop.w);
    loop.noise = moved(loop.noise);
    loop.cnt = moved(loop.cnt);
    loop.col_id = moved(loop.col_id);
    loop.holders = moved(loop.holders);
    loop.pull_begin = moved(loop.pull_begin);
    loop.peer_id = moved(loop.peer_id);
    loop.peers = &peers;
    loop.rows.chunks = moved(loop.rows.chunks);
    loop.rows.val_s = moved(loop.rows.val_s);
    loop.rows.idx_s = moved(loop.rows.idx_s);
    loop.rows.warp_begin = row_warp_begin;
    loop.cols.chunks = moved(loop.cols.chunks);
    loop.cols.val_s = moved(loop.cols.val_s);
    loop.cols.idx_s = moved(loop.cols.idx_s);
    loop.cols.warp_begin = col_warp_begin;

    const double total = header->total;
    const uint32_t n_cols_present = header->n_cols_present;
    const uint32_t * col_src = header->col_src;
    const bool all_shared = header->all_shared != 0;

    // every CTA of the group is running and has its slice before anyone touches a peer's shared memory
    if (MULTI) { cluster.sync(); } else { __syncthreads(); }

#ifdef RPVG_B200_EM_TIMELINE
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(timeline_t1));
#endif

    // ---- EM iterations. When everything this CTA touches sits in shared memory the loop runs on pointers the compiler
    // can prove to be shared (plain LDS/STS); otherwise on generic pointers (hybrid residency).
    uint32_t it;
    const double * final_a = loop.a;

    if (all_shared) {

        auto rebase = [&](auto * ptr) {
            using T = std::remove_pointer_t<decltype(ptr)>;
            if (ptr == nullptr) { return static_cast<T *>(nullptr); }
            const uint32_t off = static_cast<uint32_t>(reinterpret_cast<const unsigned char *>(ptr) - smem_raw);
            return reinterpret_cast<T *>(smem_raw + off);
        };

        SellLoop<IndexT> fast = loop;
        fast.a = rebase(loop.a);
        fast.a_prev = rebase(loop.a_prev);
        fast.g = rebase(loop.g);
        fast.w = rebase(loop.w);
        fast.noise = rebase(loop.noise);
        fast.cnt = rebase(loop.cnt);
        fast.col_id = rebase(loop.col_id);
        fast.holders = rebase(loop.holders);
        fast.pull_begin = rebase(loop.pull_begin);
        fast.peer_id = rebase(loop.peer_id);
        fast.rows.chunks = rebase(loop.rows.chunks);
        fast.rows.val_s = rebase(loop.rows.val_s);
        fast.rows.idx_s = rebase(loop.rows.idx_s);
        fast.cols.chunks = rebase(loop.cols.chunks);
        fast.cols.val_s = rebase(loop.cols.val_s);
        fast.cols.idx_s = rebase(loop.cols.idx_s);

        if constexpr (MULTI) {
            it = emSellIterateGroup<THREADS, true, IndexT>(fast, args.params, S, rank, total, red, &final_a);
        } else {
            it = emSellIterateSingle<THREADS, true, IndexT>(fast, args.params, N, total, red);
        }

    } else {

        if constexpr (MULTI) {
            it = emSellIterateGroup<THREADS, false, IndexT>(loop, args.params, S, rank, total, red, &final_a);
        } else {
            it = emSellIterateSingle<THREADS, false, IndexT>(loop, args.params, N, total, red);
        }
    }

    const double min_abundance = args.params.min_em_abundance;
    const uint16_t * holders = loop.holders;

    // rpvg:src/path_abundance_estimator.cpp:100-113. In a group every column is written by the first CTA that holds it
    // (all holders have the same value).
    double noise_part = 0;

    if (MULTI) {

        for (uint32_t l = tid; l < n_cols_present; l += THREADS) {

            if ((__ffs(holders[l]) - 1) != static_cast<int>(rank)) { continue; }

            const uint32_t j = col_src[l];
            const double aj = final_a[l];
            if (aj < min_abundance) {
                noise_part += aj * total;
                args.abundances[p0 + j] = 0;
            } else {
                args.abundances[p0 + j] = aj * total;
            }
        }

    } else {

        for (uint32_t j = tid; j < N; j += THREADS) {
            const double aj = final_a[j];
            if (aj < min_abundance) {
                noise_part += aj * total;
                args.abundances[p0 + j] = 0;
            } else {
                args.abundances[p0 + j] = aj * total;
            }
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

        // every CTA holds the noise column
        args.noise_count[c] = low_total + final_a[MULTI ? n_cols_present : N] * total;
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
            rec[0] = THREADS; rec[1] = S; rec[2] = rank; rec[3] = c; rec[4] = smid; rec[5] = n_cols_present; rec[6] = N;
            rec[7] = timeline_t0; rec[8] = timeline_t1; rec[9] = timeline_t2; rec[10] = it; rec[11] = all_shared ? 1u : 0u;
        }
    }
#endif

    // no CTA may exit while a peer can still touch its shared memory
    if (MULTI) { cluster.sync(); }
}
