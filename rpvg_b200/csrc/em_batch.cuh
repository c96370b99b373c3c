// The device-resident batch behind rpvg_b200_batch_* (em.cu) and the pipelines that feed the EM from the device
// (subset.cu): one cudaMallocAsync region with the arena, the sorted / column-major copies, results and the EM scratch.
#pragma once

#include <vector>

#include "common.cuh"

struct DevicePool {

    // One cudaMallocAsync region, carved in 256-byte aligned pieces.
    char * base = nullptr;
    size_t size = 0;
    size_t used = 0;

    template <typename T>
    T * take(size_t n) {
        used = (used + 255) & ~size_t(255);
        T * p = reinterpret_cast<T *>(base + used);
        used += n * sizeof(T);
        return p;
    }
};

template <typename T>
inline size_t padded(size_t n) { return ((n * sizeof(T)) + 255) & ~size_t(255); }

struct rpvg_b200_batch {

    uint32_t n_clusters = 0;
    uint64_t n_reads = 0;
    uint64_t n_paths = 0;
    uint64_t nnz = 0;

    std::vector<uint64_t> h_read_offset;
    std::vector<uint64_t> h_path_offset;
    std::vector<uint64_t> h_cluster_nnz;   // path non-zeros per cluster
    std::vector<uint32_t> h_iterations;
    std::vector<uint32_t> h_order;         // clusters in launch order (bins are ranges of it)

    // launch plan of the EM: clusters binned by thread-block-cluster size, shared-memory tier and index width
    struct EmBin {
        uint32_t group_size;   // CTAs per path cluster; 0 = fallback kernel (one CTA, everything in global memory)
        uint32_t threads;
        uint32_t smem;
        uint32_t index_bytes;  // 2 or 4
        uint32_t first;        // offset into d_order / d_group_scratch / d_group_cta_bytes
        uint32_t count;
        uint32_t dense_dim;    // > 0: emDenseKernel of this side (one CTA per cluster, no scratch)
        bool stream;           // emStreamKernel: one cooperative launch of the whole GPU per cluster (em_stream.cuh)
    };
    std::vector<EmBin> bins;
    uint64_t scratch_bytes = 0;

    DevicePool pool;

    // inputs
    uint64_t * d_read_offset = nullptr;
    uint64_t * d_path_offset = nullptr;
    uint64_t * d_row_ptr = nullptr;
    uint32_t * d_col = nullptr;
    double * d_prob = nullptr;
    double * d_noise = nullptr;
    uint32_t * d_count = nullptr;
    uint32_t * d_order = nullptr;
    uint64_t * d_group_scratch = nullptr;
    uint64_t * d_group_cta_bytes = nullptr;

    // built: read classes re-ordered by first path (sorted CSR), then the column-major copy
    uint64_t * d_rowkey_in = nullptr;
    uint64_t * d_rowkey_out = nullptr;
    uint32_t * d_rowid_in = nullptr;
    uint32_t * d_perm = nullptr;            // sorted position -> original read class
    uint64_t * d_sorted_len = nullptr;
    uint64_t * d_sorted_row_ptr = nullptr;
    uint32_t * d_sorted_col = nullptr;
    double * d_sorted_noise = nullptr;
    void * d_rowsort_temp = nullptr;
    size_t rowsort_temp_bytes = 0;
    void * d_scan_temp = nullptr;
    size_t scan_temp_bytes = 0;
    int rowsort_bits = 64;
    double * d_val = nullptr;
    double * d_cnt = nullptr;
    uint32_t * d_entry_row = nullptr;
    uint32_t * d_key_in = nullptr;
    uint32_t * d_key_out = nullptr;
    uint32_t * d_idx_in = nullptr;
    uint32_t * d_idx_out = nullptr;
    uint64_t * d_col_ptr = nullptr;
    uint32_t * d_csc_row = nullptr;
    double * d_csc_val = nullptr;
    void * d_sort_temp = nullptr;
    size_t sort_temp_bytes = 0;
    int sort_bits = 32;

    // EM state + results
    double * d_abund = nullptr;   // fallback kernel only: per cluster N_c + 1 relative abundances (noise last)
    double * d_w = nullptr;       // fallback kernel only
    unsigned char * d_scratch = nullptr;
    double * d_stream_partial = nullptr;      // emStreamKernel: [2 x kStreamMaxCtas] doubles + [kStreamMaxCtas] votes
    uint32_t * d_stream_flags = nullptr;
    double * d_abundances = nullptr;
    double * d_noise_count = nullptr;
    double * d_total_count = nullptr;
    uint32_t * d_iterations = nullptr;
    uint32_t * d_error = nullptr;

    bool have_results = false;

    // rows come from a device pipeline: values are already normalised (noise column included) and counts are doubles
    bool prenormalised = false;
    double * d_count_f64 = nullptr;

    // device timestamps around the two phases of rpvg_b200_batch_em (bench.py's roofline numbers)
    cudaEvent_t ev_start = nullptr;
    cudaEvent_t ev_built = nullptr;
    cudaEvent_t ev_done = nullptr;
};


// Plans the EM launches for clusters described by host-side shapes, allocates the batch and uploads offsets and plan.
// The five data arrays (d_row_ptr, d_col, d_prob, d_noise and d_count or d_count_f64) are left for the caller to fill
// (host-to-device for an uploaded arena, device-to-device for a pipeline). cluster_nnz[c] = entries of cluster c.
int rpvgBatchCreate(rpvg_b200_ctx * ctx, uint32_t n_clusters, const uint64_t * read_offset, const uint64_t * path_offset, const uint64_t * cluster_nnz,
                    bool prenormalised, rpvg_b200_batch ** out_batch);
