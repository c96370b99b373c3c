// rpvg_b200_subset_em — the "many small EMs" of the haplotype-transcripts model (one per sampled haplotype combination)
// and the single restricted EM of the strains model, as one batched device pipeline:
//
//   per subset s of the paths of cluster c(s):
//     constructPartialProbabilityMatrix   rpvg:src/path_estimator.cpp:79-113    columns restricted to the subset
//     addNoiseAndNormalizeProbabilityMatrix                   :156-166          rows re-normalised over the kept columns
//     readCollapseProbabilityMatrix                           :219-259          collapse.cuh
//     EMAbundanceEstimator                rpvg:src/path_abundance_estimator.cpp:47-114   em_sell.cuh
//   which is the body of the loop at rpvg:src/path_abundance_estimator.cpp:625-671 (Nested) and :260-269 (Minimum).
//
// Every (subset, read class) pair is a row of a device-side arena; nothing goes back to the host between the stages
// except the class counts and entry totals the EM launch plan needs (one small copy).

#include <vector>

#include <cub/device/device_scan.cuh>

#include "collapse.cuh"
#include "em_batch.cuh"

int rpvgCollapseDevice(rpvg_b200_ctx * ctx, CollapseRows rows, uint32_t * d_row_problem, double prob_precision, uint32_t * d_sorted, uint32_t * d_sorted_class,
                       uint32_t * d_n_classes);

namespace {

struct SubsetArgs {

    uint32_t n_subsets;
    uint64_t n_rows;                       // sum over subsets of the read classes of their cluster
    const uint32_t * subset_cluster;       // [n_subsets]
    const uint64_t * subset_offset;        // [n_subsets + 1] into subset_paths
    const uint64_t * subset_row_offset;    // [n_subsets + 1] rows of each subset problem
    const uint64_t * subset_map_offset;    // [n_subsets + 1] into path_map (one slot per path of the subset's cluster)
    const uint64_t * read_offset;          // arena
    const uint64_t * path_offset;
    const uint64_t * row_ptr;
    const uint32_t * col;
    const double * prob;
    const double * noise;
    const uint32_t * count;
    int32_t * path_map;                    // cluster path -> index in the subset, -1 = dropped
};

__global__ void subsetMapKernel(const SubsetArgs a, const uint32_t * __restrict__ subset_paths, uint64_t n_subset_paths, uint32_t * __restrict__ error) {

    const uint64_t k = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (k >= n_subset_paths) { return; }

    const uint32_t s = clusterOf(a.subset_offset, a.n_subsets, k);
    const uint32_t c = a.subset_cluster[s];
    const uint64_t num_paths = a.path_offset[c + 1] - a.path_offset[c];
    const uint32_t path = subset_paths[k];

    const bool ascending = (k == a.subset_offset[s]) || subset_paths[k - 1] < path;

    if (path >= num_paths || !ascending) {
        atomicOr(error, RPVG_DEV_ERR_COL_RANGE);
        return;
    }

    a.path_map[a.subset_map_offset[s] + path] = static_cast<int32_t>(k - a.subset_offset[s]);
}

// One thread per (subset, read class): number of entries that survive the restriction.
__global__ void subsetCountKernel(const SubsetArgs a, uint64_t * __restrict__ len) {

    const uint64_t q = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (q > a.n_rows) { return; }

    uint64_t kept = 0;

    if (q < a.n_rows) {

        const uint32_t s = clusterOf(a.subset_row_offset, a.n_subsets, q);
        const uint32_t c = a.subset_cluster[s];
        const uint64_t r = a.read_offset[c] + (q - a.subset_row_offset[s]);
        const int32_t * map = a.path_map + a.subset_map_offset[s];
        const uint64_t num_paths = a.path_offset[c + 1] - a.path_offset[c];

        for (uint64_t e = a.row_ptr[r]; e < a.row_ptr[r + 1]; ++e) {
            const uint32_t j = a.col[e];
            kept += (j < num_paths && map[j] >= 0) ? 1 : 0;
        }
    }

    len[q] = kept;
}

// One thread per (subset, read class): restricted, re-normalised row (exact operation order of the reference).
__global__ void subsetFillKernel(const SubsetArgs a, const uint64_t * __restrict__ out_ptr, uint32_t * __restrict__ out_col, double * __restrict__ out_val,
                                 double * __restrict__ out_noise, double * __restrict__ out_cnt) {

    const uint64_t q = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (q >= a.n_rows) { return; }

    const uint32_t s = clusterOf(a.subset_row_offset, a.n_subsets, q);
    const uint32_t c = a.subset_cluster[s];
    const uint64_t r = a.read_offset[c] + (q - a.subset_row_offset[s]);
    const int32_t * map = a.path_map + a.subset_map_offset[s];
    const uint64_t num_paths = a.path_offset[c + 1] - a.path_offset[c];

    double row_sum = 0;
    for (uint64_t e = a.row_ptr[r]; e < a.row_ptr[r + 1]; ++e) {
        const uint32_t j = a.col[e];
        if (j < num_paths && map[j] >= 0) { row_sum += a.prob[e]; }
    }

    const double nz = a.noise[r];
    const double scale = 1 - nz;
    uint64_t o = out_ptr[q];

    for (uint64_t e = a.row_ptr[r]; e < a.row_ptr[r + 1]; ++e) {

        const uint32_t j = a.col[e];
        if (j < num_paths && map[j] >= 0) {
            double v = (a.prob[e] / row_sum) * scale;
            if (isnan(v)) { v = 0; }
            out_col[o] = static_cast<uint32_t>(map[j]);
            out_val[o] = v;
            ++o;
        }
    }

    out_noise[q] = nz;
    out_cnt[q] = static_cast<double>(a.count[r]);
}

// per class: entries of its representative (one extra zero so that the scan yields n_classes + 1 offsets)
__global__ void classLengthKernel(uint64_t n_classes, const uint32_t * __restrict__ class_rep, const uint64_t * __restrict__ row_ptr, uint64_t * __restrict__ len) {

    const uint64_t g = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (g > n_classes) { return; }
    len[g] = (g < n_classes) ? row_ptr[class_rep[g] + 1] - row_ptr[class_rep[g]] : 0;
}

__global__ void gatherBoundsKernel(uint32_t n, const uint64_t * __restrict__ class_offset, const uint64_t * __restrict__ class_ptr, uint64_t * __restrict__ bounds) {

    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p <= n) { bounds[p] = class_ptr[class_offset[p]]; }
}

// per class: copy the representative's row into the EM arena
__global__ void compactClassesKernel(uint64_t n_classes, const uint32_t * __restrict__ class_rep, const double * __restrict__ class_count,
                                     const uint64_t * __restrict__ row_ptr, const uint32_t * __restrict__ col, const double * __restrict__ val,
                                     const double * __restrict__ noise, const uint64_t * __restrict__ class_ptr, uint32_t * __restrict__ out_col,
                                     double * __restrict__ out_val, double * __restrict__ out_noise, double * __restrict__ out_cnt) {

    const uint64_t g = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (g >= n_classes) { return; }

    const uint32_t rep = class_rep[g];
    uint64_t o = class_ptr[g];

    for (uint64_t e = row_ptr[rep]; e < row_ptr[rep + 1]; ++e, ++o) {
        out_col[o] = col[e];
        out_val[o] = val[e];
    }

    out_noise[g] = noise[rep];
    out_cnt[g] = class_count[g];
}

}

int rpvgBatchGibbs(rpvg_b200_ctx * ctx, rpvg_b200_batch * b, const uint32_t * num_samples, uint32_t thin_its, const uint64_t * seeds, double * abundance_samples,
                   double * noise_samples);

extern "C" int rpvg_b200_subset_em_gibbs(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, const rpvg_b200_subsets * ss, const rpvg_b200_em_params * params,
                                         double prob_precision, const rpvg_b200_gibbs * gibbs, double * abundances, double * noise_count, double * total_count,
                                         uint32_t * iterations, uint32_t * n_classes) {

    if (!ctx || !cl || !ss || !params) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (ss->n_subsets == 0 || cl->n_clusters == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no subsets / no clusters"); }
    { const int shape = rpvgCheckClusters(ctx, cl); if (shape != RPVG_B200_OK) { return shape; } }

    const uint32_t S = ss->n_subsets;
    const uint32_t C = cl->n_clusters;
    const uint64_t R = cl->read_offset[C];
    const uint64_t Z = R > 0 ? cl->row_ptr[R] : 0;
    const uint64_t SP = ss->subset_offset[S];

    // host-side shapes of the subset problems
    std::vector<uint64_t> subset_row_offset(S + 1, 0), subset_map_offset(S + 1, 0), subset_path_offset(S + 1, 0);
    uint64_t entry_bound = 0;

    for (uint32_t s = 0; s < S; ++s) {

        const uint32_t c = ss->subset_cluster[s];
        if (c >= C) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "subset_cluster out of range"); }
        if (ss->subset_offset[s + 1] <= ss->subset_offset[s]) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "empty path subset"); }

        subset_row_offset[s + 1] = subset_row_offset[s] + (cl->read_offset[c + 1] - cl->read_offset[c]);
        subset_map_offset[s + 1] = subset_map_offset[s] + (cl->path_offset[c + 1] - cl->path_offset[c]);
        subset_path_offset[s + 1] = ss->subset_offset[s + 1];
        entry_bound += cl->row_ptr[cl->read_offset[c + 1]] - cl->row_ptr[cl->read_offset[c]];
    }

    const uint64_t RS = subset_row_offset[S];
    const uint64_t MP = subset_map_offset[S];

    if (RS >= 0xffffffffull || entry_bound >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "subset batch too large: split it (rows and entries below 2^32)"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;

    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, static_cast<uint64_t *>(nullptr), static_cast<uint64_t *>(nullptr), static_cast<int>(RS + 1), st);

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    size_t bytes = 2 * pad(8 * (C + 1)) + pad(8 * (R + 1)) + pad(4 * Z) + pad(8 * Z) + pad(8 * R) + pad(4 * R);         // arena
    bytes += pad(4 * S) + 5 * pad(8 * (S + 1)) + pad(4 * SP) + pad(4 * MP);                                             // subsets + map
    bytes += 3 * pad(8 * (RS + 1)) + pad(4 * entry_bound) + pad(8 * entry_bound) + 2 * pad(8 * RS);                     // restricted rows
    bytes += 5 * pad(4 * RS) + pad(4 * S) + pad(8 * RS) + pad(scan_bytes) + pad(4) + 4096;                             // collapse + classes

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, st));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_read_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_path_offset = reinterpret_cast<uint64_t *>(take(8 * (C + 1)));
    uint64_t * d_row_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_col = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_prob = reinterpret_cast<double *>(take(8 * Z));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_count = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_subset_cluster = reinterpret_cast<uint32_t *>(take(4 * S));
    uint64_t * d_subset_offset = reinterpret_cast<uint64_t *>(take(8 * (S + 1)));
    uint64_t * d_subset_row_offset = reinterpret_cast<uint64_t *>(take(8 * (S + 1)));
    uint64_t * d_subset_map_offset = reinterpret_cast<uint64_t *>(take(8 * (S + 1)));
    uint64_t * d_class_offset = reinterpret_cast<uint64_t *>(take(8 * (S + 1)));
    uint64_t * d_bounds = reinterpret_cast<uint64_t *>(take(8 * (S + 1)));
    uint32_t * d_subset_paths = reinterpret_cast<uint32_t *>(take(4 * SP));
    int32_t * d_path_map = reinterpret_cast<int32_t *>(take(4 * MP));
    uint64_t * d_len = reinterpret_cast<uint64_t *>(take(8 * (RS + 1)));
    uint64_t * d_out_ptr = reinterpret_cast<uint64_t *>(take(8 * (RS + 1)));
    uint64_t * d_class_ptr = reinterpret_cast<uint64_t *>(take(8 * (RS + 1)));
    uint32_t * d_out_col = reinterpret_cast<uint32_t *>(take(4 * entry_bound));
    double * d_out_val = reinterpret_cast<double *>(take(8 * entry_bound));
    double * d_out_noise = reinterpret_cast<double *>(take(8 * RS));
    double * d_out_cnt = reinterpret_cast<double *>(take(8 * RS));
    uint32_t * d_row_problem = reinterpret_cast<uint32_t *>(take(4 * RS));
    uint32_t * d_sorted = reinterpret_cast<uint32_t *>(take(4 * RS));
    uint32_t * d_sorted_class = reinterpret_cast<uint32_t *>(take(4 * RS));
    uint32_t * d_class_rep = reinterpret_cast<uint32_t *>(take(4 * RS));
    uint32_t * d_unused_row_class = reinterpret_cast<uint32_t *>(take(4 * RS));
    uint32_t * d_n_classes = reinterpret_cast<uint32_t *>(take(4 * S));
    double * d_class_count = reinterpret_cast<double *>(take(8 * RS));
    void * d_scan_temp = take(scan_bytes);
    uint32_t * d_error = reinterpret_cast<uint32_t *>(take(4));
    (void) d_unused_row_class;

    int status = RPVG_B200_OK;
    cudaError_t err = cudaSuccess;
    rpvg_b200_batch * batch = nullptr;

    #define RPVG_STEP(call) if (status == RPVG_B200_OK && err == cudaSuccess) { err = (call); }
    #define RPVG_H2D(dst, src, n) if (status == RPVG_B200_OK && err == cudaSuccess && (n) > 0) { err = cudaMemcpyAsync((dst), (src), (n), cudaMemcpyHostToDevice, st); }

    RPVG_H2D(d_read_offset, cl->read_offset, 8 * (C + 1))
    RPVG_H2D(d_path_offset, cl->path_offset, 8 * (C + 1))
    RPVG_H2D(d_row_ptr, cl->row_ptr, 8 * (R + 1))
    RPVG_H2D(d_col, cl->col, 4 * Z)
    RPVG_H2D(d_prob, cl->prob, 8 * Z)
    RPVG_H2D(d_noise, cl->noise, 8 * R)
    RPVG_H2D(d_count, cl->count, 4 * R)
    RPVG_H2D(d_subset_cluster, ss->subset_cluster, 4 * S)
    RPVG_H2D(d_subset_offset, ss->subset_offset, 8 * (S + 1))
    RPVG_H2D(d_subset_row_offset, subset_row_offset.data(), 8 * (S + 1))
    RPVG_H2D(d_subset_map_offset, subset_map_offset.data(), 8 * (S + 1))
    RPVG_H2D(d_subset_paths, ss->subset_paths, 4 * SP)
    RPVG_STEP(cudaMemsetAsync(d_path_map, 0xff, 4 * MP, st))
    RPVG_STEP(cudaMemsetAsync(d_error, 0, 4, st))

    SubsetArgs a;
    a.n_subsets = S;
    a.n_rows = RS;
    a.subset_cluster = d_subset_cluster;
    a.subset_offset = d_subset_offset;
    a.subset_row_offset = d_subset_row_offset;
    a.subset_map_offset = d_subset_map_offset;
    a.read_offset = d_read_offset;
    a.path_offset = d_path_offset;
    a.row_ptr = d_row_ptr;
    a.col = d_col;
    a.prob = d_prob;
    a.noise = d_noise;
    a.count = d_count;
    a.path_map = d_path_map;

    std::vector<uint32_t> h_n_classes(S, 0);
    std::vector<uint64_t> class_offset(S + 1, 0), bounds(S + 1, 0);

    if (status == RPVG_B200_OK && err == cudaSuccess) {

        subsetMapKernel<<<static_cast<unsigned>((SP + 255) / 256), 256, 0, st>>>(a, d_subset_paths, SP, d_error);
        subsetCountKernel<<<static_cast<unsigned>((RS + 1 + 255) / 256), 256, 0, st>>>(a, d_len);
        ctx->launches += 2;
        err = cudaGetLastError();

        size_t temp = scan_bytes;
        RPVG_STEP(cub::DeviceScan::ExclusiveSum(d_scan_temp, temp, d_len, d_out_ptr, static_cast<int>(RS + 1), st))
        ctx->launches += 2;

        if (err == cudaSuccess && RS > 0) {
            subsetFillKernel<<<static_cast<unsigned>((RS + 255) / 256), 256, 0, st>>>(a, d_out_ptr, d_out_col, d_out_val, d_out_noise, d_out_cnt);
            ctx->launches++;
            err = cudaGetLastError();
        }
    }

    CollapseRows rows;
    rows.n_problems = S;
    rows.n_rows = RS;
    rows.problem_row_offset = d_subset_row_offset;
    rows.row_problem = d_row_problem;
    rows.row_ptr = d_out_ptr;
    rows.col = d_out_col;
    rows.val = d_out_val;
    rows.noise = d_out_noise;
    rows.cnt = d_out_cnt;

    if (status == RPVG_B200_OK && err == cudaSuccess) {
        RPVG_STEP(cudaMemsetAsync(d_n_classes, 0, 4 * S, st))
        if (err == cudaSuccess) { status = rpvgCollapseDevice(ctx, rows, d_row_problem, prob_precision, d_sorted, d_sorted_class, d_n_classes); }
    }

    uint32_t dev_error = 0;
    RPVG_STEP(cudaMemcpyAsync(h_n_classes.data(), d_n_classes, 4 * S, cudaMemcpyDeviceToHost, st))
    RPVG_STEP(cudaMemcpyAsync(&dev_error, d_error, 4, cudaMemcpyDeviceToHost, st))
    RPVG_STEP(cudaStreamSynchronize(st))

    if (status == RPVG_B200_OK && err == cudaSuccess && dev_error) {
        status = rpvgFail(ctx, RPVG_B200_ERR_INVALID, "malformed subsets: path index out of range or subset_paths not strictly ascending");
    }

    uint64_t n_total = 0;

    if (status == RPVG_B200_OK && err == cudaSuccess) {

        for (uint32_t s = 0; s < S; ++s) { class_offset[s + 1] = class_offset[s] + h_n_classes[s]; }
        n_total = class_offset[S];

        RPVG_H2D(d_class_offset, class_offset.data(), 8 * (S + 1))

        if (err == cudaSuccess && RS > 0) {
            collapseClassesKernel<<<static_cast<unsigned>((RS + 255) / 256), 256, 0, st>>>(rows, d_sorted, d_sorted_class, d_class_offset, nullptr, d_class_rep, d_class_count);
            ctx->launches++;
            err = cudaGetLastError();
        }

        if (err == cudaSuccess) {
            classLengthKernel<<<static_cast<unsigned>((n_total + 1 + 255) / 256), 256, 0, st>>>(n_total, d_class_rep, d_out_ptr, d_len);
            ctx->launches++;
            err = cudaGetLastError();
        }

        size_t temp = scan_bytes;
        RPVG_STEP(cub::DeviceScan::ExclusiveSum(d_scan_temp, temp, d_len, d_class_ptr, static_cast<int>(n_total + 1), st))

        if (err == cudaSuccess) {
            gatherBoundsKernel<<<(S + 1 + 255) / 256, 256, 0, st>>>(S, d_class_offset, d_class_ptr, d_bounds);
            ctx->launches += 3;
            err = cudaGetLastError();
        }

        RPVG_STEP(cudaMemcpyAsync(bounds.data(), d_bounds, 8 * (S + 1), cudaMemcpyDeviceToHost, st))
        RPVG_STEP(cudaStreamSynchronize(st))
    }

    // the collapsed classes become the read classes of an EM arena whose clusters are the subsets
    if (status == RPVG_B200_OK && err == cudaSuccess) {

        std::vector<uint64_t> cluster_nnz(S);
        for (uint32_t s = 0; s < S; ++s) { cluster_nnz[s] = bounds[s + 1] - bounds[s]; }

        status = rpvgBatchCreate(ctx, S, class_offset.data(), subset_path_offset.data(), cluster_nnz.data(), true, &batch);
    }

    if (status == RPVG_B200_OK && err == cudaSuccess) {

        RPVG_STEP(cudaMemcpyAsync(batch->d_row_ptr, d_class_ptr, 8 * (n_total + 1), cudaMemcpyDeviceToDevice, st))

        if (err == cudaSuccess && n_total > 0) {
            compactClassesKernel<<<static_cast<unsigned>((n_total + 255) / 256), 256, 0, st>>>(n_total, d_class_rep, d_class_count, d_out_ptr, d_out_col, d_out_val, d_out_noise,
                                                                                               d_class_ptr, batch->d_col, batch->d_prob, batch->d_noise, batch->d_count_f64);
            ctx->launches++;
            err = cudaGetLastError();
        }
    }

    #undef RPVG_STEP
    #undef RPVG_H2D

    if (status == RPVG_B200_OK && err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_subset_em: ") + cudaGetErrorString(err)); }

    if (status == RPVG_B200_OK) { status = rpvg_b200_batch_em(ctx, batch, params); }
    if (status == RPVG_B200_OK) { status = rpvg_b200_batch_download(ctx, batch, abundances, noise_count, total_count, iterations); }

    // rpvg:src/path_abundance_estimator.cpp:675-697: the sampler runs on the collapsed matrix of the subset
    if (status == RPVG_B200_OK && gibbs) {
        if (!gibbs->num_samples || !gibbs->seeds || !gibbs->abundance_samples || !gibbs->noise_samples) {
            status = rpvgFail(ctx, RPVG_B200_ERR_INVALID, "gibbs: NULL array");
        } else {
            status = rpvgBatchGibbs(ctx, batch, gibbs->num_samples, gibbs->thin_its, gibbs->seeds, gibbs->abundance_samples, gibbs->noise_samples);
        }
    }

    if (status == RPVG_B200_OK && n_classes) {
        for (uint32_t s = 0; s < S; ++s) { n_classes[s] = h_n_classes[s]; }
    }

    if (batch) { rpvg_b200_batch_free(ctx, batch); }
    cudaFreeAsync(base, st);

    return status;
}

extern "C" int rpvg_b200_subset_em(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * cl, const rpvg_b200_subsets * ss, const rpvg_b200_em_params * params,
                                   double prob_precision, double * abundances, double * noise_count, double * total_count, uint32_t * iterations,
                                   uint32_t * n_classes) {

    return rpvg_b200_subset_em_gibbs(ctx, cl, ss, params, prob_precision, nullptr, abundances, noise_count, total_count, iterations, n_classes);
}
