// Sort + merge of a cluster's read records before they reach the estimators — the step main.cpp runs per cluster just
// before estimate(), for all clusters of a batch at once.
//
// Reference semantics restated (never its code): rpvg:src/main.cpp:953-973 sorts the cluster's ReadPathProbabilities with
// operator< (rpvg:src/read_path_probabilities.cpp:283-322: noise probability, number of probability groups, then per
// group probability / number of paths / path ids, finally read count; two doubles are "equal" under Utils::doubleCompare,
// relative 100 eps) and then scans once: a record joins the LAST KEPT record iff quickMergeIdentical says so (:223-250:
// noise and every group probability differ by less than prob_precision, same path ids), adding its read count to it
// (uint32 arithmetic); the kept record keeps its own values.
//
// Device side: the same three steps as the read-class collapse of the matrices (collapse.cuh): bitonic network per
// cluster with the reference's comparator (data-oblivious, so the tolerance comparator cannot break it; ties keep input
// order), the greedy scan cut into independent stretches, kept records and counts. Exact: the partition, the order of the
// kept records and their counts. A record is a ragged list of lists, so this is pointer chasing, not streaming: the
// bound is L2/HBM latency; the work is O(R log^2 R) comparisons per cluster.

#include <string>

#include "collapse.cuh"
#include "em_batch.cuh"

namespace {

struct ReadRecords {

    uint32_t n_clusters;
    uint64_t n_records;
    const uint64_t * read_offset;   // [n_clusters + 1]
    const uint64_t * group_ptr;     // [n_records + 1]
    const double * group_prob;
    const uint64_t * id_ptr;        // [n_groups + 1]
    const uint32_t * ids;
    const double * noise;
    const uint32_t * count;
};

struct RecordLess {

    ReadRecords rec;
    uint64_t base;   // first record of the cluster: keys are cluster-local

    __device__ bool operator()(const uint32_t ka, const uint32_t kb) const {

        const uint64_t a = base + ka, b = base + kb;

        if (!collapseDoubleCompare(rec.noise[a], rec.noise[b])) { return rec.noise[a] < rec.noise[b]; }

        const uint64_t ga = rec.group_ptr[a], gb = rec.group_ptr[b];
        const uint64_t na = rec.group_ptr[a + 1] - ga, nb = rec.group_ptr[b + 1] - gb;
        if (na != nb) { return na < nb; }

        for (uint64_t i = 0; i < na; ++i) {

            const double pa = rec.group_prob[ga + i], pb = rec.group_prob[gb + i];
            if (!collapseDoubleCompare(pa, pb)) { return pa < pb; }

            const uint64_t ia = rec.id_ptr[ga + i], ib = rec.id_ptr[gb + i];
            const uint64_t sa = rec.id_ptr[ga + i + 1] - ia, sb = rec.id_ptr[gb + i + 1] - ib;
            if (sa != sb) { return sa < sb; }

            for (uint64_t j = 0; j < sa; ++j) {
                const uint32_t xa = rec.ids[ia + j], xb = rec.ids[ib + j];
                if (xa != xb) { return xa < xb; }
            }
        }

        if (rec.count[a] != rec.count[b]) { return rec.count[a] < rec.count[b]; }
        return false;
    }
};

// quickMergeIdentical's predicate: may record b be folded into record a?
__device__ bool recordsMergeable(const ReadRecords & rec, const uint64_t a, const uint64_t b, const double prob_precision) {

    if (fabs(rec.noise[a] - rec.noise[b]) >= prob_precision) { return false; }

    const uint64_t ga = rec.group_ptr[a], gb = rec.group_ptr[b];
    const uint64_t na = rec.group_ptr[a + 1] - ga;
    if (na != rec.group_ptr[b + 1] - gb) { return false; }

    for (uint64_t i = 0; i < na; ++i) {

        if (fabs(rec.group_prob[ga + i] - rec.group_prob[gb + i]) >= prob_precision) { return false; }

        const uint64_t ia = rec.id_ptr[ga + i], ib = rec.id_ptr[gb + i];
        const uint64_t sa = rec.id_ptr[ga + i + 1] - ia;
        if (sa != rec.id_ptr[gb + i + 1] - ib) { return false; }

        for (uint64_t j = 0; j < sa; ++j) {
            if (rec.ids[ia + j] != rec.ids[ib + j]) { return false; }
        }
    }

    return true;
}

__global__ void recordIotaKernel(const ReadRecords rec, uint32_t * __restrict__ sorted) {

    const uint32_t c = blockIdx.x;
    const uint64_t r0 = rec.read_offset[c];
    const uint32_t n = static_cast<uint32_t>(rec.read_offset[c + 1] - r0);
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) { sorted[r0 + i] = i; }
}

__global__ void __launch_bounds__(kCollapseSortThreads) recordSortKernel(const ReadRecords rec, uint32_t * __restrict__ sorted) {

    const uint32_t c = blockIdx.x;
    const uint64_t r0 = rec.read_offset[c];
    const uint32_t n = static_cast<uint32_t>(rec.read_offset[c + 1] - r0);

    RecordLess less;
    less.rec = rec;
    less.base = r0;

    collapseBitonicNetwork(sorted + r0, n, less);
}

// The scan of main.cpp:959-972 in three parallel passes, like the read-class collapse (collapse.cuh): a record that
// certainly cannot be folded into whatever record is kept before it starts a kept record of its own whatever happened
// earlier — it differs from its predecessor in structure (groups, path ids: a kept record and everything folded into it
// share them), or by at least twice the precision in the noise or a group probability (triangle inequality against the
// predecessor's kept record). Pass 1 flags these certain starts, pass 2 walks every stretch between two of them with the
// exact sequential rule, pass 3 numbers the kept records and adds up their counts.

// true iff record b certainly cannot be folded into a record that record a was folded into (or into a itself)
__device__ bool recordsCertainlyApart(const ReadRecords & rec, const uint64_t a, const uint64_t b, const double distance) {

    if (fabs(rec.noise[a] - rec.noise[b]) >= distance) { return true; }

    const uint64_t ga = rec.group_ptr[a], gb = rec.group_ptr[b];
    const uint64_t na = rec.group_ptr[a + 1] - ga;
    if (na != rec.group_ptr[b + 1] - gb) { return true; }

    for (uint64_t i = 0; i < na; ++i) {

        if (fabs(rec.group_prob[ga + i] - rec.group_prob[gb + i]) >= distance) { return true; }

        const uint64_t ia = rec.id_ptr[ga + i], ib = rec.id_ptr[gb + i];
        const uint64_t sa = rec.id_ptr[ga + i + 1] - ia;
        if (sa != rec.id_ptr[gb + i + 1] - ib) { return true; }

        for (uint64_t j = 0; j < sa; ++j) {
            if (rec.ids[ia + j] != rec.ids[ib + j]) { return true; }
        }
    }

    return false;
}

// pass 1: one thread per sorted position
__global__ void recordBoundaryKernel(const ReadRecords rec, const uint32_t * __restrict__ sorted, const double prob_precision, uint32_t * __restrict__ flag) {

    const uint64_t s = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (s >= rec.n_records) { return; }

    const uint32_t c = clusterOf(rec.read_offset, rec.n_clusters, s);
    const uint64_t r0 = rec.read_offset[c];

    bool start = (s == r0);
    if (!start) { start = recordsCertainlyApart(rec, r0 + sorted[s - 1], r0 + sorted[s], 2.0 * prob_precision * (1.0 + 1e-9)); }

    flag[s] = start ? 1u : 0u;
}

// pass 2: the thread of a certain start walks its stretch with the sequential rule; records kept inside it get flag 2
__global__ void recordStretchKernel(const ReadRecords rec, const uint32_t * __restrict__ sorted, const double prob_precision, uint32_t * flag) {

    const uint64_t s = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (s >= rec.n_records || flag[s] != 1u) { return; }

    const uint32_t c = clusterOf(rec.read_offset, rec.n_clusters, s);
    const uint64_t r0 = rec.read_offset[c];
    const uint64_t r1 = rec.read_offset[c + 1];

    uint64_t rep = r0 + sorted[s];

    for (uint64_t t = s + 1; t < r1; ++t) {

        if (flag[t] == 1u) { break; }

        const uint64_t r = r0 + sorted[t];
        if (!recordsMergeable(rec, rep, r, prob_precision)) {
            rep = r;
            flag[t] = 2u;
        }
    }
}

constexpr int kRecordIndexThreads = 256;

// pass 3, one CTA per cluster: index of the kept record of every sorted position (block scan of the start flags), the kept
// records themselves and — by the thread of each start — the sum of the counts folded into it (uint32 arithmetic, like
// ReadPathProbabilities::addReadCount: exact whatever the order)
__global__ void __launch_bounds__(kRecordIndexThreads) recordIndexKernel(const ReadRecords rec, const uint32_t * __restrict__ sorted, const uint32_t * __restrict__ flag,
                                                                         uint32_t * __restrict__ record_class, uint32_t * __restrict__ n_merged,
                                                                         uint32_t * __restrict__ merged_rep, uint32_t * __restrict__ merged_count) {

    __shared__ uint32_t warp_total[kRecordIndexThreads / 32];
    __shared__ uint32_t carry;

    const uint32_t c = blockIdx.x;
    const uint64_t r0 = rec.read_offset[c];
    const uint64_t r1 = rec.read_offset[c + 1];

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) { carry = 0; }
    __syncthreads();

    for (uint64_t base = r0; base < r1; base += kRecordIndexThreads) {

        const uint64_t s = base + tid;
        const uint32_t start = (s < r1 && flag[s] != 0u) ? 1u : 0u;

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

        if (s < r1) {

            const uint32_t kept = before + incl - 1;
            record_class[r0 + sorted[s]] = kept;

            if (start) {

                uint32_t total = rec.count[r0 + sorted[s]];
                for (uint64_t t = s + 1; t < r1 && flag[t] == 0u; ++t) { total += rec.count[r0 + sorted[t]]; }

                merged_rep[r0 + kept] = sorted[s];
                merged_count[r0 + kept] = total;
            }
        }

        __syncthreads();
        if (tid == kRecordIndexThreads - 1) { carry = before + incl; }
        __syncthreads();
    }

    if (tid == 0) { n_merged[c] = carry; }
}

}

extern "C" int rpvg_b200_merge_read_records(rpvg_b200_ctx * ctx, const rpvg_b200_read_records * in, double prob_precision, uint32_t * n_merged, uint32_t * record_class,
                                            uint32_t * merged_rep, uint32_t * merged_count) {

    if (!ctx || !in) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_merge_read_records: NULL argument"); }
    if (in->n_clusters == 0) { return RPVG_B200_OK; }
    if (!in->read_offset || !n_merged) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_merge_read_records: NULL array"); }

    const uint32_t C = in->n_clusters;
    const uint64_t R = in->read_offset[C];

    if (R == 0) {
        for (uint32_t c = 0; c < C; ++c) { n_merged[c] = 0; }
        return RPVG_B200_OK;
    }

    if (!in->group_ptr || !in->noise || !in->count || !record_class || !merged_rep || !merged_count) {
        return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_merge_read_records: NULL array");
    }

    const uint64_t G = in->group_ptr[R];
    const uint64_t I = G > 0 ? in->id_ptr[G] : 0;

    uint64_t largest = 0;
    for (uint32_t c = 0; c < C; ++c) {
        if (in->read_offset[c + 1] < in->read_offset[c]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_merge_read_records: read_offset must be non-decreasing"); }
        largest = std::max<uint64_t>(largest, in->read_offset[c + 1] - in->read_offset[c]);
    }
    if (largest >= (1ull << 31) || R >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_merge_read_records: too many records"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    const size_t bytes = padded<uint64_t>(C + 1) + padded<uint64_t>(R + 1) + padded<double>(G + 1) + padded<uint64_t>(G + 1) + padded<uint32_t>(I + 1) + padded<double>(R) +
                         padded<uint32_t>(R) + 5 * padded<uint32_t>(R) + padded<uint32_t>(C) + 4096;

    DevicePool pool;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&pool.base), bytes, ctx->pool, s));
    const DeviceAllocationGuard allocation_guard{s, pool.base};
    pool.size = bytes;

    uint64_t * d_read_offset = pool.take<uint64_t>(C + 1);
    uint64_t * d_group_ptr = pool.take<uint64_t>(R + 1);
    double * d_group_prob = pool.take<double>(G + 1);
    uint64_t * d_id_ptr = pool.take<uint64_t>(G + 1);
    uint32_t * d_ids = pool.take<uint32_t>(I + 1);
    double * d_noise = pool.take<double>(R);
    uint32_t * d_count = pool.take<uint32_t>(R);
    uint32_t * d_sorted = pool.take<uint32_t>(R);
    uint32_t * d_class = pool.take<uint32_t>(R);
    uint32_t * d_rep = pool.take<uint32_t>(R);
    uint32_t * d_merged_count = pool.take<uint32_t>(R);
    uint32_t * d_n_merged = pool.take<uint32_t>(C);
    uint32_t * d_flag = pool.take<uint32_t>(R);

    cudaError_t err = cudaSuccess;
    auto up = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); }
    };

    up(d_read_offset, in->read_offset, (C + 1) * sizeof(uint64_t));
    up(d_group_ptr, in->group_ptr, (R + 1) * sizeof(uint64_t));
    up(d_group_prob, in->group_prob, G * sizeof(double));
    up(d_id_ptr, in->id_ptr, G > 0 ? (G + 1) * sizeof(uint64_t) : 0);
    up(d_ids, in->ids, I * sizeof(uint32_t));
    up(d_noise, in->noise, R * sizeof(double));
    up(d_count, in->count, R * sizeof(uint32_t));

    if (err == cudaSuccess) { err = cudaMemsetAsync(d_rep, 0, R * sizeof(uint32_t), s); }
    if (err == cudaSuccess) { err = cudaMemsetAsync(d_merged_count, 0, R * sizeof(uint32_t), s); }

    if (err == cudaSuccess) {

        ReadRecords rec;
        rec.n_clusters = C;
        rec.n_records = R;
        rec.read_offset = d_read_offset;
        rec.group_ptr = d_group_ptr;
        rec.group_prob = d_group_prob;
        rec.id_ptr = d_id_ptr;
        rec.ids = d_ids;
        rec.noise = d_noise;
        rec.count = d_count;

        recordIotaKernel<<<C, 256, 0, s>>>(rec, d_sorted);
        RPVG_LAUNCHED(ctx);
        recordSortKernel<<<C, kCollapseSortThreads, 0, s>>>(rec, d_sorted);
        RPVG_LAUNCHED(ctx);
        const unsigned rgrid = static_cast<unsigned>((R + 255) / 256);
        recordBoundaryKernel<<<rgrid, 256, 0, s>>>(rec, d_sorted, prob_precision, d_flag);
        RPVG_LAUNCHED(ctx);
        recordStretchKernel<<<rgrid, 256, 0, s>>>(rec, d_sorted, prob_precision, d_flag);
        RPVG_LAUNCHED(ctx);
        recordIndexKernel<<<C, kRecordIndexThreads, 0, s>>>(rec, d_sorted, d_flag, d_class, d_n_merged, d_rep, d_merged_count);
        RPVG_LAUNCHED(ctx);

        err = cudaGetLastError();
    }

    auto down = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, s); }
    };

    down(n_merged, d_n_merged, C * sizeof(uint32_t));
    down(record_class, d_class, R * sizeof(uint32_t));
    down(merged_rep, d_rep, R * sizeof(uint32_t));
    down(merged_count, d_merged_count, R * sizeof(uint32_t));

    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }


    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_merge_read_records: ") + cudaGetErrorString(err)); }
    return RPVG_B200_OK;
}
