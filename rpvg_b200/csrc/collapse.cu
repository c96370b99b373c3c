// rpvg_b200_collapse — read-class collapse of a batch of (already normalised) probability matrices. See collapse.cuh.

#include <vector>

#include "collapse.cuh"

namespace {

__global__ void collapseInitKernel(uint32_t n_problems, uint64_t n_rows, const uint64_t * __restrict__ problem_row_offset, uint32_t * __restrict__ row_problem,
                                   uint32_t * __restrict__ sorted) {

    const uint64_t r = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (r >= n_rows) { return; }

    row_problem[r] = clusterOf(problem_row_offset, n_problems, r);
    sorted[r] = static_cast<uint32_t>(r);
}

}

// Sort + scan on rows that are already on the device (nothing is allocated here). d_sorted [n_rows], d_sorted_class [n_rows], d_n_classes [n_problems],
// d_row_problem [n_rows] are outputs.
int rpvgCollapseDevice(rpvg_b200_ctx * ctx, CollapseRows rows, uint32_t * d_row_problem, double prob_precision, uint32_t * d_sorted, uint32_t * d_sorted_class,
                       uint32_t * d_n_classes) {

    cudaStream_t s = ctx->stream;
    if (rows.n_rows == 0) { return RPVG_B200_OK; }

    const unsigned grid = static_cast<unsigned>((rows.n_rows + 255) / 256);
    collapseInitKernel<<<grid, 256, 0, s>>>(rows.n_problems, rows.n_rows, rows.problem_row_offset, d_row_problem, d_sorted);
    RPVG_LAUNCHED(ctx);

    rows.row_problem = d_row_problem;

    collapseSortKernel<<<rows.n_problems, kCollapseSortThreads, 0, s>>>(rows, d_sorted);
    RPVG_LAUNCHED(ctx);

    collapseBoundaryKernel<<<grid, 256, 0, s>>>(rows, d_sorted, prob_precision, d_sorted_class);
    RPVG_LAUNCHED(ctx);
    collapseStretchKernel<<<static_cast<unsigned>((rows.n_rows + kCollapseStretchThreads - 1) / kCollapseStretchThreads), kCollapseStretchThreads, 0, s>>>(rows, d_sorted, prob_precision,
                                                                                                                                                        d_sorted_class);
    RPVG_LAUNCHED(ctx);
    collapseIndexKernel<<<rows.n_problems, kCollapseIndexThreads, 0, s>>>(rows, d_sorted_class, d_n_classes);
    RPVG_LAUNCHED(ctx);

    return RPVG_B200_OK;
}

extern "C" int rpvg_b200_collapse(rpvg_b200_ctx * ctx, const rpvg_b200_collapse_problems * pr, double prob_precision, uint32_t * row_class, uint32_t * n_classes,
                                  uint32_t * class_rep, double * class_count) {

    if (!ctx || !pr || !n_classes) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "NULL argument"); }
    if (pr->n_problems == 0) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "no problems"); }

    const uint32_t P = pr->n_problems;
    const uint64_t R = pr->row_offset[P];
    const uint64_t Z = R > 0 ? pr->row_ptr[R] : 0;

    if (R >= 0xffffffffull || Z >= 0xffffffffull) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "too many rows or entries in one batch"); }

    if (R == 0) {
        for (uint32_t p = 0; p < P; ++p) { n_classes[p] = 0; }
        return RPVG_B200_OK;
    }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    auto pad = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bytes = 2 * pad(8 * (P + 1)) + pad(8 * (R + 1)) + pad(4 * Z) + pad(8 * Z) + 2 * pad(8 * R) + 5 * pad(4 * R) + pad(4 * P) + pad(8 * R) + 1024;

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));

    size_t used = 0;
    auto take = [&](size_t b) { char * q = base + used; used += pad(b); return q; };

    uint64_t * d_row_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_class_offset = reinterpret_cast<uint64_t *>(take(8 * (P + 1)));
    uint64_t * d_row_ptr = reinterpret_cast<uint64_t *>(take(8 * (R + 1)));
    uint32_t * d_col = reinterpret_cast<uint32_t *>(take(4 * Z));
    double * d_val = reinterpret_cast<double *>(take(8 * Z));
    double * d_noise = reinterpret_cast<double *>(take(8 * R));
    double * d_cnt = reinterpret_cast<double *>(take(8 * R));
    uint32_t * d_row_problem = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_sorted = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_sorted_class = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_row_class = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_class_rep = reinterpret_cast<uint32_t *>(take(4 * R));
    uint32_t * d_n_classes = reinterpret_cast<uint32_t *>(take(4 * P));
    double * d_class_count = reinterpret_cast<double *>(take(8 * R));

    int status = RPVG_B200_OK;
    cudaError_t err = cudaSuccess;

    #define RPVG_H2D(dst, src, n)                                                                                  \
        if (err == cudaSuccess && (n) > 0) { err = cudaMemcpyAsync((dst), (src), (n), cudaMemcpyHostToDevice, s); }

    RPVG_H2D(d_row_offset, pr->row_offset, 8 * (P + 1))
    RPVG_H2D(d_row_ptr, pr->row_ptr, 8 * (R + 1))
    RPVG_H2D(d_col, pr->col, 4 * Z)
    RPVG_H2D(d_val, pr->val, 8 * Z)
    RPVG_H2D(d_noise, pr->noise, 8 * R)
    RPVG_H2D(d_cnt, pr->count, 8 * R)
    #undef RPVG_H2D

    if (err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("collapse H2D: ") + cudaGetErrorString(err)); }

    CollapseRows rows;
    rows.n_problems = P;
    rows.n_rows = R;
    rows.problem_row_offset = d_row_offset;
    rows.row_problem = d_row_problem;
    rows.row_ptr = d_row_ptr;
    rows.col = d_col;
    rows.val = d_val;
    rows.noise = d_noise;
    rows.cnt = d_cnt;

    if (status == RPVG_B200_OK) { status = rpvgCollapseDevice(ctx, rows, d_row_problem, prob_precision, d_sorted, d_sorted_class, d_n_classes); }

    std::vector<uint64_t> class_offset(P + 1, 0);

    if (status == RPVG_B200_OK) {

        err = cudaMemcpyAsync(n_classes, d_n_classes, 4 * P, cudaMemcpyDeviceToHost, s);
        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

        if (err == cudaSuccess) {
            for (uint32_t p = 0; p < P; ++p) { class_offset[p + 1] = class_offset[p] + n_classes[p]; }
            err = cudaMemcpyAsync(d_class_offset, class_offset.data(), 8 * (P + 1), cudaMemcpyHostToDevice, s);
        }

        if (err == cudaSuccess) {
            const unsigned grid = static_cast<unsigned>((R + 255) / 256);
            collapseClassesKernel<<<grid, 256, 0, s>>>(rows, d_sorted, d_sorted_class, d_class_offset, d_row_class, d_class_rep, d_class_count);
            ctx->launches++;
            err = cudaGetLastError();
        }

        const uint64_t n_total = class_offset[P];
        if (err == cudaSuccess && row_class) { err = cudaMemcpyAsync(row_class, d_row_class, 4 * R, cudaMemcpyDeviceToHost, s); }
        if (err == cudaSuccess && class_rep) { err = cudaMemcpyAsync(class_rep, d_class_rep, 4 * n_total, cudaMemcpyDeviceToHost, s); }
        if (err == cudaSuccess && class_count) { err = cudaMemcpyAsync(class_count, d_class_count, 8 * n_total, cudaMemcpyDeviceToHost, s); }
        if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }

        if (err != cudaSuccess) { status = rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("collapse: ") + cudaGetErrorString(err)); }
    }

    cudaFreeAsync(base, s);
    return status;
}
