// rpvg_b200_grouped_collapse — the matrix the haplotype-transcripts model computes its haplotype posteriors on.
//
// Reference semantics restated (never its code), for every cluster:
//   constructGroupedProbabilityMatrix   rpvg:src/path_estimator.cpp:115-154   column g = sum of the probabilities of the
//                                                                              paths of group g (a path may sit in
//                                                                              several groups)
//   addNoiseAndNormalizeProbabilityMatrix                  :156-166
//   readCollapseProbabilityMatrix                          :219-259          collapse.cuh
//   noise column split off again         rpvg:src/path_abundance_estimator.cpp:445-446
// With one group per path of a transcript this is also constructPartialProbabilityMatrix + normalise + collapse of
// inferAbundancesIndependentGroups (rpvg:src/path_abundance_estimator.cpp:377-383).
//
// The grouped matrix is dense (as in the reference): one thread per read class accumulates its row, so the additions
// into a cell happen in a fixed order (ascending path index; the reference adds in probability order, the difference is
// an ulp and far inside the collapse's own tolerance).

#include <vector>

#include <cub/device/device_scan.cuh>

#include "collapse.cuh"
#include "grouped.cuh"

int rpvgCollapseDevice(rpvg_b200_ctx * ctx, CollapseRows rows, uint32_t * d_row_problem, double prob_precision, uint32_t * d_sorted, uint32_t * d_sorted_class,
                       uint32_t * d_n_classes);

namespace {

struct GroupedArgs {

    uint32_t n_clusters;
    uint64_t n_reads;
    const uint64_t * read_offset;
    const uint64_t * path_offset;
    const uint64_t * row_ptr;
    const uint32_t * col;
    const double * prob;
    const double * noise;
    const uint32_t * count;
    const uint64_t * group_offset;       // [n_clusters + 1]
    const uint64_t * matrix_offset;      // [n_clusters + 1] dense block of cluster c: rows_c x groups_c, column-major
    const uint64_t * path_group_ptr;     // [n_paths + 1] groups of every (global) path: inverse of the group lists
    const uint32_t * path_groups;        // cluster-local group indices
    double * matrix;
};

// one thread per read class: grouped row, normalised in place; number of non-zeros for the sparse copy
__global__ void groupedRowsKernel(const GroupedArgs a, uint64_t * __restrict__ nnz_len) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r > a.n_reads) { return; }
    if (r == a.n_reads) { nnz_len[r] = 0; return; }

    const uint32_t c = clusterOf(a.read_offset, a.n_clusters, r);
    const uint64_t rows = a.read_offset[c + 1] - a.read_offset[c];
    const uint64_t groups = a.group_offset[c + 1] - a.group_offset[c];
    const uint64_t local = r - a.read_offset[c];
    const uint64_t path_base = a.path_offset[c];
    const uint64_t num_paths = a.path_offset[c + 1] - path_base;

    double * row = a.matrix + a.matrix_offset[c] + local;   // element g at row[g * rows]

    // accumulate (the block is zero-filled)
    for (uint64_t e = a.row_ptr[r]; e < a.row_ptr[r + 1]; ++e) {

        const uint32_t j = a.col[e];
        if (j >= num_paths) { continue; }

        const double p = a.prob[e];
        for (uint64_t k = a.path_group_ptr[path_base + j]; k < a.path_group_ptr[path_base + j + 1]; ++k) { row[a.path_groups[k] * rows] += p; }
    }

    // normalise: left-to-right row sum, (x / sum) * (1 - noise), NaN -> 0
    double row_sum = 0;
    for (uint64_t g = 0; g < groups; ++g) { row_sum += row[g * rows]; }

    const double scale = 1 - a.noise[r];
    uint64_t kept = 0;

    for (uint64_t g = 0; g < groups; ++g) {
        double v = (row[g * rows] / row_sum) * scale;
        if (isnan(v)) { v = 0; }
        row[g * rows] = v;
        kept += (v != 0);
    }

    nnz_len[r] = kept;
}

// dense rows -> CSR (non-zeros) for the collapse
__global__ void groupedSparseKernel(const GroupedArgs a, const uint64_t * __restrict__ out_ptr, uint32_t * __restrict__ out_col, double * __restrict__ out_val,
                                    double * __restrict__ out_cnt) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= a.n_reads) { return; }

    const uint32_t c = clusterOf(a.read_offset, a.n_clusters, r);
    const uint64_t rows = a.read_offset[c + 1] - a.read_offset[c];
    const uint64_t groups = a.group_offset[c + 1] - a.group_offset[c];
    const double * row = a.matrix + a.matrix_offset[c] + (r - a.read_offset[c]);

    uint64_t o = out_ptr[r];
    for (uint64_t g = 0; g < groups; ++g) {
        const double v = row[g * rows];
        if (v != 0) {
            out_col[o] = static_cast<uint32_t>(g);
            out_val[o] = v;
            ++o;
        }
    }

    out_cnt[r] = static_cast<double>(a.count[r]);
}

// one thread per class: the representative's dense row, compacted (leading dimension = classes of the cluster)
__global__ void groupedGatherKernel(const GroupedArgs a, uint64_t n_classes_total, const uint64_t * __restrict__ class_offset, const uint32_t * __restrict__ class_rep,
                                    const double * __restrict__ class_count, double * __restrict__ out_matrix, double * __restrict__ out_noise,
                                    double * __restrict__ out_counts) {

    const uint64_t g_idx = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (g_idx >= n_classes_total) { return; }

    const uint32_t rep = class_rep[g_idx];
    const uint32_t c = clusterOf(a.read_offset, a.n_clusters, rep);
    const uint64_t rows = a.read_offset[c + 1] - a.read_offset[c];
    const uint64_t groups = a.group_offset[c + 1] - a.group_offset[c];
    const uint64_t classes = class_offset[c + 1] - class_offset[c];
    const uint64_t k = g_idx - class_offset[c];

    const double * row = a.matrix + a.matrix_offset[c] + (rep - a.read_offset[c]);
    double * out = out_matrix + a.matrix_offset[c] + k;

    for (uint64_t g = 0; g < groups; ++g) { out[g * classes] = row[g * rows]; }

    out_noise[a.read_offset[c] + k] = a.noise[rep];
    out_counts[a.read_offset[c] + k] = class_count[g_idx];
}

}

// keep == nullptr: results to the host arrays; otherwise they stay on the device under *keep (matrix/noise/counts unused)
static int groupedCollapse(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, const rpvg_b200_path_groups * gr, double prob_precision, uint32_t * n_classes, double * matrix,
                           double * noise, double * counts, rpvg_b200_grouped ** keep) {

    if (!ctx || !cl || !gr || !n_classes || (!keep && (!matrix || !noise || !counts))) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (cl->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no clusters"); }
    { const int shape = rpvgCheckClusters(ctx, cl); if (shape != RPVG_B200_OK) { return shape; } }

    const uint32_t C = cl->n_clusters;
    const uint64_t R = cl->read_offset[C];
    const uint64_t P = cl->path_offset[C];
    const uint64_t Z = R > 0 ? cl->row_ptr[R] : 0;
    const uint64_t G = gr->group_offset[C];
    const uint64_t GP = gr->group_path_ptr[G];

    // host: dense block offsets and the inverse index path -> groups (counting sort; group order within a path ascending,
    // which is the order the reference's path_id_group_idx is filled in)
    std::vector<uint64_t> matrix_offset(C + 1, 0), path_group_ptr(P + 1, 0);
    std::vector<uint32_t> path_groups(GP);

    for (uint32_t c = 0; c < C; ++c) {

        const uint64_t rows = cl->read_offset[c + 1] - cl->read_offset[c];
        const uint64_t groups = gr->group_offset[c + 1] - gr->group_offset[c];
        const uint64_t num_paths = cl->path_offset[c + 1] - cl->path_offset[c];

        if (groups == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "a cluster has no path groups"); }
        matrix_offset[c + 1] = matrix_offset[c] + rows * groups;

        for (uint64_t g = gr->group_offset[c]; g < gr->group_offset[c + 1]; ++g) {
            if (gr->group_path_ptr[g + 1] <= gr->group_path_ptr[g]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "empty path group"); }
            for (uint64_t k = gr->group_path_ptr[g]; k < gr->group_path_ptr[g + 1]; ++k) {
                if (gr->group_paths[k] >= num_paths) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "group path index out of range"); }
                path_group_ptr[cl->path_offset[c] + gr->group_paths[k] + 1]++;
            }
        }
    }

    for (uint64_t j = 0; j < P; ++j) { path_group_ptr[j + 1] += path_group_ptr[j]; }

    {
        std::vector<uint64_t> cursor(path_group_ptr.begin(), path_group_ptr.end() - 1);
        for (uint32_t c = 0; c < C; ++c) {
            for (uint64_t g = gr->group_offset[c]; g < gr->group_offset[c + 1]; ++g) {
                for (uint64_t k = gr->group_path_ptr[g]; k < gr->group_path_ptr[g + 1]; ++k) {
                    path_groups[cursor[cl->path_offset[c] + gr->group_paths[k]]++] = static_cast<uint32_t>(g - gr->group_offset[c]);
                }
            }
        }
    }

    const uint64_t MX = matrix_offset[C];
    if (R >= 0xffffffffull || MX >= (uint64_t(1) << 33)) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "grouped matrix too large for one call (rows x groups)"); }

    if (R == 0) {
        for (uint32_t c = 0; c < C; ++c) { n_classes[c] = 0; }
        if (keep) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_create: no read classes"); }
        return RPVG_B200_OK;
    }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr), static_cast<int>(R + 1), s);

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    size_t bytes = 5 * pad(8 * (C + 1)) + pad(8 * (R + 1)) + pad(4 * Z) + pad(8 * Z) + pad(8 * R) + pad(4 * R) + pad(8 * (P + 1)) + pad(4 * GP);
    bytes += 2 * pad(8 * MX) + 2 * pad(8 * (R + 1)) + pad(4 * MX) + pad(8 * MX) + 3 * pad(8 * R) + 4 * pad(4 * R) + pad(4 * C) + pad(scan_bytes) + 4096;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_read_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_group_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_matrix_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_class_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_row_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_col = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_prob = reinterpret_cast<double *>(take(8 * Z));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_count = reinterpret_cast<uint32_t *>(take(4 * R));
    uint64_t * d_path_group_ptr = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint32_t * d_path_groups = reinterpret_cast<uint32_t *>(take(4 * GP));
    double * d_matrix = reinterpret_cast<double *>(take(8 * MX));
    double * d_out_matrix = reinterpret_cast<double *>(take(8 * MX));
    uint64_t * d_len = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint64_t * d_out_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_out_col = reinterpret_cast<uint32_t *>(take(4 * MX));
    double * d_out_val = reinterpret_cast<double *>(take(8 * MX));
    double * d_out_cnt = reinterpret_cast<double *>(take(8 * R));
    double * d_class_count = reinterpret_cast<double *>(take(8 * R));
    double * d_res_noise = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_row_problem = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_sorted = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_sorted_class = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_class_rep = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_n_classes = reinterpret_cast<uint32_t *>(take(4 * C));
    void * d_scan_temp = take(scan_bytes);
    double * d_res_counts = reinterpret_cast<double *>(d_len);   // the length buffer is free again once the scan has run

    cudaError_t err = cudaSuccess;
    int status = RPVG_B200_OK;

    #define RPVG_STEP(call) if (status == RPVG_B200_OK && err == cudaSuccess) { err = (call); }
    #define RPVG_H2D(dst, src, n) if (status == RPVG_B200_OK && err == cudaSuccess && (n) > 0) { err = cudaMemcpyAsync((dst), (src), (n), cudaMemcpyHostToDevice, s); }

    RPVG_H2D(d_read_offset, cl->read_offset, 8 * (C + 1))
    RPVG_H2D(d_path_offset, cl->path_offset, 8 * (C + 1))
    RPVG_H2D(d_group_offset, gr->group_offset, 8 * (C + 1))
    RPVG_H2D(d_matrix_offset, matrix_offset.data(), 8 * (C + 1))
    RPVG_H2D(d_row_ptr, cl->row_ptr, 8 * (R + 1))
    RPVG_H2D(d_col, cl->col, 4 * Z)
    RPVG_H2D(d_prob, cl->prob, 8 * Z)
    RPVG_H2D(d_noise, cl->noise, 8 * R)
    RPVG_H2D(d_count, cl->count, 4 * R)
    RPVG_H2D(d_path_group_ptr, path_group_ptr.data(), 8 * (P + 1))
    RPVG_H2D(d_path_groups, path_groups.data(), 4 * GP)
    RPVG_STEP(cudaMemsetAsync(d_matrix, 0, 8 * MX, s))

    GroupedArgs a;
    a.n_clusters = C;
    a.n_reads = R;
    a.read_offset = d_read_offset;
    a.path_offset = d_path_offset;
    a.row_ptr = d_row_ptr;
    a.col = d_col;
    a.prob = d_prob;
    a.noise = d_noise;
    a.count = d_count;
    a.group_offset = d_group_offset;
    a.matrix_offset = d_matrix_offset;
    a.path_group_ptr = d_path_group_ptr;
    a.path_groups = d_path_groups;
    a.matrix = d_matrix;

    if (status == RPVG_B200_OK && err == cudaSuccess) {
        groupedRowsKernel<<<static_cast<unsigned>((R + 1 + 255) / 256), 256, 0, s>>>(a, d_len);
        ctx->launches++;
        err = cudaGetLastError();
    }

    size_t temp = scan_bytes;
    RPVG_STEP(cub::DeviceScan::ExclusiveSum(d_scan_temp, temp, d_len, d_out_ptr, static_cast<int>(R + 1), s))

    if (status == RPVG_B200_OK && err == cudaSuccess) {
        groupedSparseKernel<<<static_cast<unsigned>((R + 255) / 256), 256, 0, s>>>(a, d_out_ptr, d_out_col, d_out_val, d_out_cnt);
        ctx->launches += 3;
        err = cudaGetLastError();
    }

    CollapseRows rows;
    rows.n_problems = C;
    rows.n_rows = R;
    rows.problem_row_offset = d_read_offset;
    rows.row_problem = d_row_problem;
    rows.row_ptr = d_out_ptr;
    rows.col = d_out_col;
    rows.val = d_out_val;
    rows.noise = d_noise;
    rows.cnt = d_out_cnt;

    if (status == RPVG_B200_OK && err == cudaSuccess) { status = rpvgCollapseDevice(ctx, rows, d_row_problem, prob_precision, d_sorted, d_sorted_class, d_n_classes); }

    RPVG_STEP(cudaMemcpyAsync(n_classes, d_n_classes, 4 * C, cudaMemcpyDeviceToHost, s))
    RPVG_STEP(cudaStreamSynchronize(s))

    std::vector<uint64_t> class_offset(C + 1, 0);

    if (status == RPVG_B200_OK && err == cudaSuccess) {

        for (uint32_t c = 0; c < C; ++c) { class_offset[c + 1] = class_offset[c] + n_classes[c]; }
        const uint64_t n_total = class_offset[C];

        RPVG_H2D(d_class_offset, class_offset.data(), 8 * (C + 1))

        if (err == cudaSuccess) {
            collapseClassesKernel<<<static_cast<unsigned>((R + 255) / 256), 256, 0, s>>>(rows, d_sorted, d_sorted_class, d_class_offset, nullptr, d_class_rep, d_class_count);
            groupedGatherKernel<<<static_cast<unsigned>((n_total + 255) / 256), 256, 0, s>>>(a, n_total, d_class_offset, d_class_rep, d_class_count, d_out_matrix, d_res_noise,
                                                                                             d_res_counts);
            ctx->launches += 2;
            err = cudaGetLastError();
        }

        if (!keep) {
            RPVG_STEP(cudaMemcpyAsync(matrix, d_out_matrix, 8 * MX, cudaMemcpyDeviceToHost, s))
            RPVG_STEP(cudaMemcpyAsync(noise, d_res_noise, 8 * R, cudaMemcpyDeviceToHost, s))
            RPVG_STEP(cudaMemcpyAsync(counts, d_res_counts, 8 * R, cudaMemcpyDeviceToHost, s))
        }
        RPVG_STEP(cudaStreamSynchronize(s))
    }

    #undef RPVG_STEP
    #undef RPVG_H2D

    if (status == RPVG_B200_OK && err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_grouped_collapse: ") + cudaGetErrorString(err)); }

    if (keep && status == RPVG_B200_OK) {

        auto * grouped = new rpvg_b200_grouped();
        grouped->n_clusters = C;
        grouped->base = base;
        grouped->d_matrix = d_out_matrix;
        grouped->d_noise = d_res_noise;
        grouped->d_counts = d_res_counts;
        grouped->d_row_offset = d_read_offset;
        grouped->d_group_offset = d_group_offset;
        grouped->d_matrix_offset = d_matrix_offset;
        grouped->d_n_classes = d_n_classes;
        grouped->h_rows_total = cl->read_offset[C];
        *keep = grouped;
        return status;
    }

    cudaFreeAsync(base, s);
    return status;
}

extern "C" int rpvg_b200_grouped_collapse(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, const rpvg_b200_path_groups * gr, double prob_precision,
                                          uint32_t * n_classes, double * matrix, double * noise, double * counts) {

    return groupedCollapse(ctx, cl, gr, prob_precision, n_classes, matrix, noise, counts, nullptr);
}

extern "C" int rpvg_b200_grouped_create(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, const rpvg_b200_path_groups * gr, double prob_precision,
                                        uint32_t * n_classes, rpvg_b200_grouped ** grouped) {

    if (!grouped) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    *grouped = nullptr;
    return groupedCollapse(ctx, cl, gr, prob_precision, n_classes, nullptr, nullptr, nullptr, grouped);
}

namespace {

// one thread per read class: its probabilities scattered into the zero-filled dense block of its cluster
__global__ void rawRowsKernel(const uint32_t n_clusters, const uint64_t n_reads, const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ path_offset,
                              const uint64_t * __restrict__ matrix_offset, const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col,
                              const double * __restrict__ prob, const uint32_t * __restrict__ count, double * __restrict__ matrix, double * __restrict__ counts,
                              uint32_t * __restrict__ rows_of_cluster) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_reads) { return; }

    const uint32_t c = clusterOf(read_offset, n_clusters, r);
    const uint64_t rows = read_offset[c + 1] - read_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - path_offset[c];
    const uint64_t local = r - read_offset[c];

    double * row = matrix + matrix_offset[c] + local;

    for (uint64_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) {
        const uint32_t j = col[e];
        if (j < num_paths) { row[j * rows] = prob[e]; }
    }

    counts[r] = static_cast<double>(count[r]);
    if (local == 0) { rows_of_cluster[c] = static_cast<uint32_t>(rows); }
}

}

// The raw matrices of the haplotypes model (constructProbabilityMatrix, rpvg:src/path_estimator.cpp:55-77: probabilities
// as they are, noise apart, no normalisation, no collapse) built on the device from the arena and kept there.
extern "C" int rpvg_b200_grouped_raw(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, rpvg_b200_grouped ** out) {

    if (!ctx || !cl || !out) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    *out = nullptr;
    if (cl->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no clusters"); }
    { const int shape = rpvgCheckClusters(ctx, cl); if (shape != RPVG_B200_OK) { return shape; } }

    const uint32_t C = cl->n_clusters;
    const uint64_t R = cl->read_offset[C];
    if (R == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw: no read classes"); }
    const uint64_t Z = cl->row_ptr[R];

    std::vector<uint64_t> matrix_offset(C + 1, 0);
    for (uint32_t c = 0; c < C; ++c) {
        const uint64_t rows = cl->read_offset[c + 1] - cl->read_offset[c];
        const uint64_t num_paths = cl->path_offset[c + 1] - cl->path_offset[c];
        if (rows == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw: a cluster has no read classes"); }
        matrix_offset[c + 1] = matrix_offset[c] + rows * num_paths;
    }

    const uint64_t MX = matrix_offset[C];
    if (R >= 0xffffffffull || MX >= (uint64_t(1) << 33)) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "dense matrices too large for one call (rows x paths)"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 3 * pad(8 * (C + 1)) + pad(8 * (R + 1)) + pad(4 * Z) + pad(8 * Z) + 2 * pad(8 * R) + pad(4 * R) + pad(8 * MX) + pad(4 * C) + 4096;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_read_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_matrix_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_row_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_col = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_prob = reinterpret_cast<double *>(take(8 * Z));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    double * d_counts = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_count = reinterpret_cast<uint32_t *>(take(4 * R));
    double * d_matrix = reinterpret_cast<double *>(take(8 * MX));
    uint32_t * d_rows = reinterpret_cast<uint32_t *>(take(4 * C));

    cudaError_t err = cudaSuccess;
    auto up = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); }
    };

    up(d_read_offset, cl->read_offset, 8 * (C + 1));
    up(d_path_offset, cl->path_offset, 8 * (C + 1));
    up(d_matrix_offset, matrix_offset.data(), 8 * (C + 1));
    up(d_row_ptr, cl->row_ptr, 8 * (R + 1));
    up(d_col, cl->col, 4 * Z);
    up(d_prob, cl->prob, 8 * Z);
    up(d_noise, cl->noise, 8 * R);
    up(d_count, cl->count, 4 * R);

    if (err == cudaSuccess) { err = cudaMemsetAsync(d_matrix, 0, 8 * MX, s); }

    if (err == cudaSuccess) {
        rawRowsKernel<<<static_cast<unsigned>((R + 255) / 256), 256, 0, s>>>(C, R, d_read_offset, d_path_offset, d_matrix_offset, d_row_ptr, d_col, d_prob, d_count, d_matrix,
                                                                            d_counts, d_rows);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }   // matrix_offset and the caller's arrays may go away

    if (err != cudaSuccess) {
        cudaFreeAsync(base, s);
        return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_grouped_raw: ") + cudaGetErrorString(err));
    }

    auto * grouped = new rpvg_b200_grouped();
    grouped->n_clusters = C;
    grouped->base = base;
    grouped->d_matrix = d_matrix;
    grouped->d_noise = d_noise;
    grouped->d_counts = d_counts;
    grouped->d_row_offset = d_read_offset;
    grouped->d_group_offset = d_path_offset;
    grouped->d_matrix_offset = d_matrix_offset;
    grouped->d_n_classes = d_rows;
    grouped->h_rows_total = cl->read_offset[C];
    *out = grouped;

    return RPVG_B200_OK;
}

namespace {

// one thread per read record: the probability of each of its groups scattered to the group's paths in the zero-filled
// dense block of its cluster
__global__ void rawRecordRowsKernel(const uint32_t n_clusters, const uint64_t n_reads, const uint64_t * __restrict__ read_offset, const uint64_t * __restrict__ path_offset,
                                    const uint64_t * __restrict__ matrix_offset, const uint64_t * __restrict__ group_ptr, const double * __restrict__ group_prob,
                                    const uint64_t * __restrict__ id_ptr, const uint32_t * __restrict__ ids, const uint32_t * __restrict__ count, double * __restrict__ matrix,
                                    double * __restrict__ counts, uint32_t * __restrict__ rows_of_cluster) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_reads) { return; }

    const uint32_t c = clusterOf(read_offset, n_clusters, r);
    const uint64_t rows = read_offset[c + 1] - read_offset[c];
    const uint64_t num_paths = path_offset[c + 1] - path_offset[c];
    const uint64_t local = r - read_offset[c];

    double * row = matrix + matrix_offset[c] + local;

    for (uint64_t g = group_ptr[r]; g < group_ptr[r + 1]; ++g) {
        const double p = group_prob[g];
        for (uint64_t k = id_ptr[g]; k < id_ptr[g + 1]; ++k) {
            const uint32_t j = ids[k];
            if (j < num_paths) { row[j * rows] = p; }
        }
    }

    counts[r] = static_cast<double>(count[r]);
    if (local == 0) { rows_of_cluster[c] = static_cast<uint32_t>(rows); }
}

}

// rpvg_b200_grouped_raw from read records as ReadPathProbabilities holds them (a probability and the paths that share
// it): a third of the bytes of the expanded arena go over the host link.
extern "C" int rpvg_b200_grouped_raw_records(rpvg_b200_ctx * ctx, const rpvg_b200_read_records * rec, const uint64_t * path_offset, rpvg_b200_grouped ** out) {

    if (!ctx || !rec || !path_offset || !out) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    *out = nullptr;
    if (rec->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no clusters"); }
    if (!rec->read_offset || !rec->group_ptr || !rec->id_ptr || !rec->noise || !rec->count) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: NULL array"); }

    const uint32_t C = rec->n_clusters;
    const uint64_t R = rec->read_offset[C];
    if (R == 0 || rec->read_offset[0] != 0 || path_offset[0] != 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: no read records, or offsets that do not start at 0"); }

    const uint64_t NG = rec->group_ptr[R];
    if (rec->group_ptr[0] != 0 || rec->id_ptr[0] != 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: pointers must start at 0"); }
    for (uint64_t r = 0; r < R; ++r) {
        if (rec->group_ptr[r + 1] < rec->group_ptr[r]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: group_ptr decreases"); }
    }
    for (uint64_t g = 0; g < NG; ++g) {
        if (rec->id_ptr[g + 1] < rec->id_ptr[g]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: id_ptr decreases"); }
    }
    const uint64_t Z = rec->id_ptr[NG];
    if ((NG > 0 && !rec->group_prob) || (Z > 0 && !rec->ids)) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: NULL array"); }

    std::vector<uint64_t> matrix_offset(C + 1, 0);
    for (uint32_t c = 0; c < C; ++c) {
        if (rec->read_offset[c + 1] < rec->read_offset[c] || path_offset[c + 1] < path_offset[c]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: offsets decrease"); }
        const uint64_t rows = rec->read_offset[c + 1] - rec->read_offset[c];
        const uint64_t num_paths = path_offset[c + 1] - path_offset[c];
        if (rows == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_grouped_raw_records: a cluster has no read records"); }
        if (num_paths >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_grouped_raw_records: too many paths in a cluster"); }
        matrix_offset[c + 1] = matrix_offset[c] + rows * num_paths;
    }

    const uint64_t MX = matrix_offset[C];
    if (R >= 0xffffffffull || MX >= (uint64_t(1) << 33)) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "dense matrices too large for one call (rows x paths)"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 3 * pad(8 * (C + 1)) + pad(8 * (R + 1)) + pad(8 * NG) + pad(8 * (NG + 1)) + pad(4 * Z) + 2 * pad(8 * R) + pad(4 * R) + pad(8 * MX) + pad(4 * C) + 4096;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_read_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_matrix_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_group_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    double * d_group_prob = reinterpret_cast<double *>(take(8 * NG));
    uint64_t * d_id_ptr = reinterpret_cast<uint64_t *>(take(8 * (NG + 1)));
    uint32_t * d_ids = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    double * d_counts = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_count = reinterpret_cast<uint32_t *>(take(4 * R));
    double * d_matrix = reinterpret_cast<double *>(take(8 * MX));
    uint32_t * d_rows = reinterpret_cast<uint32_t *>(take(4 * C));

    cudaError_t err = cudaSuccess;
    auto up = [&](void * dst, const void * src, size_t n) {
        if (err == cudaSuccess && n > 0) { err = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s); }
    };

    up(d_read_offset, rec->read_offset, 8 * (C + 1));
    up(d_path_offset, path_offset, 8 * (C + 1));
    up(d_matrix_offset, matrix_offset.data(), 8 * (C + 1));
    up(d_group_ptr, rec->group_ptr, 8 * (R + 1));
    up(d_group_prob, rec->group_prob, 8 * NG);
    up(d_id_ptr, rec->id_ptr, 8 * (NG + 1));
    up(d_ids, rec->ids, 4 * Z);
    up(d_noise, rec->noise, 8 * R);
    up(d_count, rec->count, 4 * R);

    if (err == cudaSuccess) { err = cudaMemsetAsync(d_matrix, 0, 8 * MX, s); }

    if (err == cudaSuccess) {
        rawRecordRowsKernel<<<static_cast<unsigned>((R + 255) / 256), 256, 0, s>>>(C, R, d_read_offset, d_path_offset, d_matrix_offset, d_group_ptr, d_group_prob, d_id_ptr, d_ids, d_count,
                                                                                  d_matrix, d_counts, d_rows);
        ctx->launches++;
        err = cudaGetLastError();
    }

    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }   // matrix_offset and the caller's arrays may go away

    if (err != cudaSuccess) {
        cudaFreeAsync(base, s);
        return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_grouped_raw_records: ") + cudaGetErrorString(err));
    }

    auto * grouped = new rpvg_b200_grouped();
    grouped->n_clusters = C;
    grouped->base = base;
    grouped->d_matrix = d_matrix;
    grouped->d_noise = d_noise;
    grouped->d_counts = d_counts;
    grouped->d_row_offset = d_read_offset;
    grouped->d_group_offset = d_path_offset;
    grouped->d_matrix_offset = d_matrix_offset;
    grouped->d_n_classes = d_rows;
    grouped->h_rows_total = R;
    *out = grouped;

    return RPVG_B200_OK;
}

extern "C" void rpvg_b200_grouped_free(rpvg_b200_ctx * ctx, rpvg_b200_grouped * grouped) {

    if (!grouped) { return; }
    if (ctx) {
        cudaSetDevice(ctx->device);
        cudaFreeAsync(grouped->base, ctx->stream);
    }
    delete grouped;
}
