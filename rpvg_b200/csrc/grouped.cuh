// The grouped + collapsed matrices of a batch of clusters kept on the device (rpvg_b200_grouped_create), so that the
// log-likelihood rounds of the haplotype posteriors read them where they are instead of a round trip through the host.
#pragma once

#include <cstdint>
#include <vector>

struct rpvg_b200_grouped {

    uint32_t n_clusters = 0;
    char * base = nullptr;                     // the one allocation everything below lives in
    const double * d_matrix = nullptr;         // cluster c at d_matrix_offset[c]: n_classes[c] x groups_c, column-major
    const double * d_noise = nullptr;          // cluster c at d_row_offset[c], n_classes[c] values
    const double * d_counts = nullptr;
    const uint64_t * d_row_offset = nullptr;   // [n_clusters + 1] (the input read offsets)
    const uint64_t * d_group_offset = nullptr; // [n_clusters + 1]
    const uint64_t * d_matrix_offset = nullptr;
    const uint32_t * d_n_classes = nullptr;    // [n_clusters]
    uint64_t h_rows_total = 0;                 // d_row_offset[n_clusters] (host copy): extent of the per-row arrays

    // results of the last rpvg_b200_grouped_bounded_pairs on this handle (host memory, valid until the next one / free)
    std::vector<uint64_t> h_cand_offset;
    std::vector<uint32_t> h_cand_pair;
    std::vector<double> h_cand_loglik;

    // device time and log evaluations of the pair kernels of all rpvg_b200_grouped_bounded_pairs calls on this handle
    double pair_ms = 0;
    double pair_logs = 0;
};
