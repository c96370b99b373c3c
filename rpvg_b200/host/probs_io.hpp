// Reader and writer of rpvg's `--write-probs` text dump (`<prefix>_probs.txt.gz`): the boundary input of the estimators,
// one block per path cluster. With it a dump made by a real rpvg run drives the GPU path (and any CPU checker) without GBWT,
// protobuf or htslib.
//
// Format restated from ProbabilityClusterWriter::addCluster (rpvg:src/threaded_output_writer.cpp:42-95), never its code:
//   #
//   name,length,effective_length name,length,effective_length ...        (8 significant digits)
//   read_count noise_prob prob:id,id,... prob:id,...                     (one line per read class; ids are the
//   ...                                                                    positions in the path line; max(8,
//                                                                          ceil(-log10 prob_precision)) digits)
// Clusters without read classes are not written. The dump is lossy (decimal, 8 digits): a workload source, not a
// bit-exact fixture. The file is gzip (BGZF blocks are gzip members), read and written here with zlib.
#ifndef RPVG_B200_HOST_PROBS_IO_HPP
#define RPVG_B200_HOST_PROBS_IO_HPP

#include <cstdint>
#include <string>
#include <vector>

namespace rpvg_b200_host {

// All clusters of a dump, flattened into the arena layout of include/rpvg_b200.h (rpvg_b200_clusters) plus the path table.
struct ProbsDump {

    std::vector<uint64_t> read_offset;    // [n_clusters + 1]
    std::vector<uint64_t> path_offset;    // [n_clusters + 1]
    std::vector<uint64_t> row_ptr;        // [n_reads + 1]
    std::vector<uint32_t> col;            // cluster-local path index, ascending within a read class
    std::vector<double> prob;
    std::vector<double> noise;            // [n_reads]
    std::vector<uint32_t> count;          // [n_reads]

    std::vector<std::string> path_name;   // [n_paths]
    std::vector<uint32_t> path_length;
    std::vector<double> path_effective_length;

    uint32_t numClusters() const { return read_offset.empty() ? 0 : static_cast<uint32_t>(read_offset.size() - 1); }
};

// Throws std::runtime_error with the line number on malformed input.
ProbsDump readProbsDump(const std::string & filename);

// prob_precision as given to rpvg (decides the number of digits of the probabilities, like the reference's writer).
void writeProbsDump(const std::string & filename, const ProbsDump & dump, double prob_precision);

}

// C view for non-C++ callers (tests, tools). Handles are freed with rpvg_b200_probs_free; array pointers stay valid until then.
extern "C" {

struct rpvg_b200_probs_view {

    uint32_t n_clusters;
    uint64_t n_reads, n_paths, nnz;
    const uint64_t * read_offset;
    const uint64_t * path_offset;
    const uint64_t * row_ptr;
    const uint32_t * col;
    const double * prob;
    const double * noise;
    const uint32_t * count;
    const uint32_t * path_length;
    const double * path_effective_length;
    const char * path_names;   // '\n'-separated, n_paths names
};

// NULL on error (message in error[0..error_len)).
void * rpvg_b200_probs_read(const char * filename, rpvg_b200_probs_view * view, char * error, uint64_t error_len);
void rpvg_b200_probs_free(void * handle);

// 0 on success. path_names: '\n'-separated.
int rpvg_b200_probs_write(const char * filename, const rpvg_b200_probs_view * view, double prob_precision, char * error, uint64_t error_len);

}

#endif
