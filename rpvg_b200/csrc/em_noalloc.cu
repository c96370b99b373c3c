// rpvg_b200_em_cluster_noalloc — EM abundance estimation of ONE path cluster without any allocation.
//
// Who needs this: callers that own all of their device memory up front and must not (or cannot) allocate while they
// run — a host thread inside a region where the CUDA allocator is off limits (a memory-budgeted service whose pool was
// sized at start-up, a real-time loop, code running under another framework's caching allocator that forbids side
// allocations, a process close to its device-memory limit). The regular entry points (em.cu) call cudaMallocAsync for
// every batch; this one calls no cudaMalloc / cudaFree / cudaMallocAsync and no library routine that allocates.
//
// How: it is the one-CTA-per-cluster kernel of em_kernels.cuh (every array in global memory) with its work area —
// normalised values, abundance vector, column accumulators — placed in a fixed array that is part of this library's
// device image (`noalloc_scratch` below, kNoAllocScratchBytes = 4 MiB). Every GPU that loads the library carries those
// 4 MiB until the process exits. The caller passes device pointers for inputs and outputs; nothing is staged.
//
// Rules that follow from the shared, fixed array:
//   * limit: 8 * (nnz + 2 * (n_paths + 1)) bytes must fit in 4 MiB, else RPVG_B200_ERR_LIMIT and nothing is launched;
//   * all calls on one device use the same array and find in it whatever the previous call left; the kernel overwrites
//     what it reads first and zeroes its accumulators at the start of every iteration, so stale contents are harmless;
//   * two calls whose kernels may overlap on the GPU (different streams, no dependency) race on the array: serialise
//     calls per device;
//   * the first call on a device looks the array's device address up (cudaGetSymbolAddress), which may also make the
//     CUDA runtime load this library's device code on that GPU. Make that first call before starting any stream capture
//     on the device; later calls only launch a kernel on the caller's stream;
//   * the call makes `device` the calling thread's current device and leaves it so;
//   * the inputs get the same checks as the regular entry point (offsets consistent with the nnz the caller states, path
//     indices in range and strictly ascending within a read class, probabilities finite and non-negative). A failed check
//     is reported on the device: d_iterations[0] = 0xffffffff, nothing else written.
// Column sums are accumulated with atomics (there is no column-major copy to gather from, building one would need more
// memory), so results agree with the regular entry points to rounding (~1e-15 relative), not bit for bit.

#include <cstdio>
#include <mutex>

#include "common.cuh"

namespace {

constexpr size_t kNoAllocScratchBytes = size_t(4) << 20;
constexpr int kNoAllocThreads = 512;
constexpr int kMaxDevices = 64;

// The work area. In host code this name is only a handle for cudaGetSymbolAddress — never a device address.
static __device__ __align__(256) unsigned char noalloc_scratch[kNoAllocScratchBytes];

std::mutex setup_mutex;
unsigned char * scratch_address[kMaxDevices] = {};   // per device, filled by the first call on that device

void noallocMessage(char * buffer, size_t buffer_len, const char * text) {

    if (buffer && buffer_len > 0) { snprintf(buffer, buffer_len, "%s", text); }
}

__global__ void __launch_bounds__(kNoAllocThreads) emNoAllocKernel(const uint32_t M, const uint32_t N, const uint64_t host_nnz, const uint64_t * __restrict__ row_ptr,
                                                                   const uint32_t * __restrict__ col, const double * __restrict__ prob,
                                                                   const double * __restrict__ noise, const uint32_t * __restrict__ count,
                                                                   const rpvg_b200_em_params params, unsigned char * scratch, double * abundances,
                                                                   double * noise_count, double * total_count, uint32_t * iterations) {

    __shared__ double red[32];

    const int tid = threadIdx.x;
    const uint64_t e_base = row_ptr[0];
    const uint64_t nnz = row_ptr[M] - e_base;

    // The host sized the work area from the caller's nnz; the layout below follows row_ptr. If the two disagree (or
    // row_ptr runs backwards) nothing is written: every thread sees the same values, so the exit is uniform.
    if (nnz != host_nnz || row_ptr[M] < e_base) {
        if (tid == 0) { *iterations = 0xffffffffu; }
        return;
    }

    double * val = reinterpret_cast<double *>(scratch);   // [nnz] normalised probabilities
    double * a = val + nnz;                               // [N + 1] relative abundances, noise last
    double * g = a + (N + 1);                             // [N + 1] column sums of the current iteration

    if (M == 0) {   // rpvg:src/path_abundance_estimator.cpp:20-22
        for (uint32_t j = tid; j < N; j += kNoAllocThreads) { abundances[j] = 0; }
        if (tid == 0) {
            *noise_count = 0;
            *total_count = 0;
            *iterations = 0;
        }
        return;
    }

    // First a pass that only reads: the checks the regular entry point makes in buildRowsKernel (em.cu) — entry offsets
    // non-decreasing and inside [row_ptr[0], row_ptr[M]], path indices in range and strictly ascending within a read
    // class, probabilities finite and non-negative. Any failure stops the call (iterations = ~0) before the work area or
    // the outputs are touched.
    int bad = 0;

    for (uint32_t i = tid; i < M; i += kNoAllocThreads) {

        const uint64_t e0 = row_ptr[i], e1 = row_ptr[i + 1];
        if (e0 < e_base || e1 < e0 || e1 - e_base > nnz) { bad = 1; continue; }

        int64_t prev_col = -1;
        for (uint64_t e = e0; e < e1; ++e) {
            const uint32_t j = col[e];
            const double p = prob[e];
            if (j >= N || static_cast<int64_t>(j) <= prev_col || !(p >= 0) || isinf(p)) { bad = 1; }
            prev_col = j;
        }
    }

    if (__syncthreads_or(bad)) {
        if (tid == 0) { *iterations = 0xffffffffu; }
        return;
    }

    // rpvg:src/path_estimator.cpp:156-166 on the non-zeros
    double count_sum = 0;

    for (uint32_t i = tid; i < M; i += kNoAllocThreads) {

        const uint64_t e0 = row_ptr[i], e1 = row_ptr[i + 1];
        double row_sum = 0;
        for (uint64_t e = e0; e < e1; ++e) { row_sum += prob[e]; }

        const double scale = 1 - noise[i];
        for (uint64_t e = e0; e < e1; ++e) {
            double v = (prob[e] / row_sum) * scale;
            if (isnan(v)) { v = 0; }
            val[e - e_base] = v;
        }
        count_sum += static_cast<double>(count[i]);
    }

    const double total = blockSum(count_sum, red);

    const double a_init = static_cast<double>(1 / static_cast<float>(N + 1));   // rpvg:src/path_abundance_estimator.cpp:54
    for (uint32_t j = tid; j <= N; j += kNoAllocThreads) { a[j] = a_init; }
    __syncthreads();

    uint32_t conv_its = 0;
    uint32_t it = 0;

    while (it < params.max_em_its) {

        for (uint32_t j = tid; j < N; j += kNoAllocThreads) { g[j] = 0; }   // the array keeps old contents: start from zero
        __syncthreads();

        const double a_noise = a[N];
        double noise_acc = 0;

        for (uint32_t i = tid; i < M; i += kNoAllocThreads) {

            const uint64_t e0 = row_ptr[i] - e_base, e1 = row_ptr[i + 1] - e_base;
            double s = 0;
            for (uint64_t e = e0; e < e1; ++e) { s += val[e] * a[col[e + e_base]]; }
            s += noise[i] * a_noise;

            const double wi = static_cast<double>(count[i]) / s;
            for (uint64_t e = e0; e < e1; ++e) { atomicAdd(&g[col[e + e_base]], val[e] * wi); }
            noise_acc += noise[i] * wi;
        }

        const double noise_g = blockSum(noise_acc, red);   // barriers inside: g is complete afterwards

        bool ok = true;

        for (uint32_t j = tid; j <= N; j += kNoAllocThreads) {

            const double a_old = a[j];
            const double a_new = (a_old * (j == N ? noise_g : g[j])) / total;
            a[j] = a_new;
            if (a_new >= params.min_em_abundance && fabs(a_new - a_old) / a_new > params.max_rel_em_conv) { ok = false; }
        }

        ++it;

        if (__syncthreads_and(ok)) {
            if (++conv_its == params.min_em_conv_its) { break; }
        } else {
            conv_its = 0;
        }
    }

    // rpvg:src/path_abundance_estimator.cpp:100-113
    double noise_part = 0;
    for (uint32_t j = tid; j < N; j += kNoAllocThreads) {
        const double aj = a[j];
        if (aj < params.min_em_abundance) {
            noise_part += aj * total;
            abundances[j] = 0;
        } else {
            abundances[j] = aj * total;
        }
    }

    const double low_sum = blockSum(noise_part, red);

    if (tid == 0) {
        *noise_count = low_sum + a[N] * total;
        *total_count = total;
        *iterations = it;
    }
}

}

extern "C" int rpvg_b200_em_cluster_noalloc(int device, void * stream, uint32_t n_reads, uint32_t n_paths, uint64_t nnz, const uint64_t * d_row_ptr,
                                            const uint32_t * d_col, const double * d_prob, const double * d_noise, const uint32_t * d_count,
                                            const rpvg_b200_em_params * params, double * d_abundances, double * d_noise_count, double * d_total_count,
                                            uint32_t * d_iterations, char * error_message, size_t error_message_len) {

    noallocMessage(error_message, error_message_len, "");

    if (device < 0 || device >= kMaxDevices || n_paths == 0 || !params || params->min_em_conv_its == 0 || !d_row_ptr || !d_abundances || !d_noise_count ||
        !d_total_count || !d_iterations || (n_reads > 0 && (!d_noise || !d_count)) || (nnz > 0 && (!d_col || !d_prob))) {
        noallocMessage(error_message, error_message_len, "rpvg_b200_em_cluster_noalloc: invalid argument (NULL pointer, no paths, or device ordinal out of range)");
        return RPVG_B200_ERR_INVALID;
    }

    // (compare before multiplying: 8 * nnz wraps for nnz >= 2^61)
    const unsigned long long slots = kNoAllocScratchBytes / 8;
    const unsigned long long vector_slots = 2ull * (static_cast<unsigned long long>(n_paths) + 1);
    const unsigned long long needed = (nnz > slots) ? ~0ull : 8ull * (nnz + vector_slots);

    if (nnz > slots || needed > kNoAllocScratchBytes) {
        char text[256];
        snprintf(text, sizeof(text), "rpvg_b200_em_cluster_noalloc: cluster needs %llu bytes of work area (8 * (nnz + 2 * (n_paths + 1))), the fixed area holds %zu; "
                                     "use rpvg_b200_em_abundances for clusters this large", needed, kNoAllocScratchBytes);
        noallocMessage(error_message, error_message_len, text);
        return RPVG_B200_ERR_LIMIT;
    }

    cudaError_t err = cudaSetDevice(device);
    if (err != cudaSuccess) {
        noallocMessage(error_message, error_message_len, cudaGetErrorString(err));
        return RPVG_B200_ERR_CUDA;
    }

    unsigned char * scratch = nullptr;
    {
        // set-up, once per device, with that device current: the symbol's address on THIS device
        std::lock_guard<std::mutex> lock(setup_mutex);

        if (!scratch_address[device]) {
            void * address = nullptr;
            err = cudaGetSymbolAddress(&address, noalloc_scratch);
            if (err != cudaSuccess || !address) {
                noallocMessage(error_message, error_message_len, err != cudaSuccess ? cudaGetErrorString(err) : "cudaGetSymbolAddress returned NULL");
                return RPVG_B200_ERR_CUDA;
            }
            scratch_address[device] = static_cast<unsigned char *>(address);
        }
        scratch = scratch_address[device];
    }

    emNoAllocKernel<<<1, kNoAllocThreads, 0, static_cast<cudaStream_t>(stream)>>>(n_reads, n_paths, nnz, d_row_ptr, d_col, d_prob, d_noise, d_count, *params, scratch,
                                                                                  d_abundances, d_noise_count, d_total_count, d_iterations);
    err = cudaGetLastError();
    if (err != cudaSuccess) {
        noallocMessage(error_message, error_message_len, cudaGetErrorString(err));
        return RPVG_B200_ERR_CUDA;
    }

    return RPVG_B200_OK;
}
