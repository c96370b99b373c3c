// Transcripts-per-million from read counts — the reducer the abundance writer applies to the estimates.
//
// Reference semantics restated (never its code): rpvg:src/main.cpp:1029-1057 sums abundance / effective_length over every
// path of every cluster (paths with effective_length <= 0 contribute nothing) into total_transcript_count;
// AbundanceEstimatesWriter::addEstimates (rpvg:src/threaded_output_writer.cpp:305-320) prints, per path,
// (abundance / effective_length) / total_transcript_count * 10^6, and 0 for a path without effective length.
//
// Device side: one pass that forms the per-path transcript counts, a cub::DeviceReduce sum (fixed tree for a given size:
// run-to-run deterministic; the reference adds in thread/cluster order, so agreement is to rounding, ~1e-15 relative),
// one pass that scales. HBM-bound streaming, 8 + 8 B read and 8 B written per path and pass.

#include <cub/device/device_reduce.cuh>

#include "common.cuh"

namespace {

__global__ void transcriptCountKernel(const uint64_t n, const double * __restrict__ abundances, const double * __restrict__ effective_lengths, double * __restrict__ transcript_count) {

    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n) { return; }

    const double len = effective_lengths[i];
    transcript_count[i] = len > 0 ? abundances[i] / len : 0.0;
}

__global__ void tpmScaleKernel(const uint64_t n, const double * __restrict__ total, double * __restrict__ transcript_count) {

    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n) { return; }

    // the reference's order of operations: count / total * 10^6
    transcript_count[i] = transcript_count[i] / *total * 1e6;
}

}

extern "C" int rpvg_b200_tpm(rpvg_b200_ctx * ctx, uint64_t n_paths, const double * abundances, const double * effective_lengths, double * tpm, double * total_transcript_count) {

    if (!ctx) { return RPVG_B200_ERR_INVALID; }
    if (n_paths == 0) {
        if (total_transcript_count) { *total_transcript_count = 0; }
        return RPVG_B200_OK;
    }
    if (!abundances || !effective_lengths || !tpm) { return rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_tpm: NULL array"); }
    if (n_paths >= (1ull << 40)) { return rpvgFail(ctx, RPVG_B200_ERR_LIMIT, "rpvg_b200_tpm: too many paths"); }

    RPVG_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    size_t reduce_bytes = 0;
    RPVG_CUDA(ctx, cub::DeviceReduce::Sum(nullptr, reduce_bytes, static_cast<const double *>(nullptr), static_cast<double *>(nullptr), static_cast<int64_t>(n_paths), s));

    const size_t array_bytes = ((n_paths * sizeof(double)) + 255) & ~size_t(255);
    const size_t bytes = 3 * array_bytes + 256 + ((reduce_bytes + 255) & ~size_t(255));

    char * base = nullptr;
    RPVG_CUDA(ctx, cudaMallocFromPoolAsync(reinterpret_cast<void **>(&base), bytes, ctx->pool, s));
    const DeviceAllocationGuard allocation_guard{s, base};

    double * d_abundances = reinterpret_cast<double *>(base);
    double * d_lengths = reinterpret_cast<double *>(base + array_bytes);
    double * d_count = reinterpret_cast<double *>(base + 2 * array_bytes);
    double * d_total = reinterpret_cast<double *>(base + 3 * array_bytes);
    void * d_reduce = base + 3 * array_bytes + 256;

    cudaError_t err = cudaMemcpyAsync(d_abundances, abundances, n_paths * sizeof(double), cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) { err = cudaMemcpyAsync(d_lengths, effective_lengths, n_paths * sizeof(double), cudaMemcpyHostToDevice, s); }

    const unsigned grid = static_cast<unsigned>((n_paths + 255) / 256);

    if (err == cudaSuccess) {
        transcriptCountKernel<<<grid, 256, 0, s>>>(n_paths, d_abundances, d_lengths, d_count);
        RPVG_LAUNCHED(ctx);
        err = cub::DeviceReduce::Sum(d_reduce, reduce_bytes, d_count, d_total, static_cast<int64_t>(n_paths), s);
        RPVG_LAUNCHED(ctx);
    }

    double total = 0;
    if (err == cudaSuccess) {
        tpmScaleKernel<<<grid, 256, 0, s>>>(n_paths, d_total, d_count);
        RPVG_LAUNCHED(ctx);
        err = cudaMemcpyAsync(tpm, d_count, n_paths * sizeof(double), cudaMemcpyDeviceToHost, s);
    }
    if (err == cudaSuccess) { err = cudaMemcpyAsync(&total, d_total, sizeof(double), cudaMemcpyDeviceToHost, s); }
    if (err == cudaSuccess) { err = cudaStreamSynchronize(s); }
    if (err == cudaSuccess) { err = cudaGetLastError(); }


    if (err != cudaSuccess) { return rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("rpvg_b200_tpm: ") + cudaGetErrorString(err)); }

    if (total_transcript_count) { *total_transcript_count = total; }
    return RPVG_B200_OK;
}
