/* Reference-only oracles around the hot path (test infrastructure, not product): rpvg's own
 * ReadPathProbabilities::addPathProbs (src/read_path_probabilities.cpp:74-221) and its own output writers
 * (src/threaded_output_writer.cpp), compiled unmodified by oracle/Makefile and driven from flat arrays. */
#ifndef RPVG_B200_TESTS_REFERENCE_IO_HARNESS_H
#define RPVG_B200_TESTS_REFERENCE_IO_HARNESS_H

#include <stdint.h>

#include "estimator_harness.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The alignment paths of the reads of ONE cluster. Read r owns alignment paths [align_ptr[r], align_ptr[r+1]), the last
 * of them the noise alignment path (no ids); alignment path a was found on the paths ids[id_ptr[a] .. id_ptr[a+1])
 * (cluster-local indices: the harness maps path id i to index i). */
typedef struct {
    uint32_t num_paths;
    const double * effective_length;   /* [num_paths] */
    const uint32_t * path_group;       /* [num_paths] group (transcript) of each path when collapse_groups, else NULL */
    const uint32_t * source_count;     /* [num_paths] or NULL (-> 1) */
    uint32_t num_groups;
    uint32_t num_reads;
    const uint8_t * min_mapq;          /* [num_reads] */
    const uint64_t * align_ptr;        /* [num_reads + 1] */
    const int32_t * score_sum;         /* [n_align] */
    const uint32_t * align_length;     /* [n_align] (uint16 in rpvg) */
    const uint32_t * frag_length;      /* [n_align] (uint16 in rpvg) */
    const uint64_t * id_ptr;           /* [n_align + 1] */
    const uint32_t * ids;
} HarnessAlignments;

/* FragmentLengthDist(mean, sd, sd_max_multi).logProb(value) */
double ref_frag_log_prob(double mean, double sd, uint32_t sd_max_multi, uint32_t value);

/* addPathProbs on every read; outputs: noise[num_reads], group_ptr[num_reads + 1], group_prob, group_id_ptr, group_ids
 * (capacity: number of input ids + 1 each). Returns the number of probability groups written. */
uint64_t ref_add_path_probs(const HarnessAlignments * al, double frag_mean, double frag_sd, uint32_t sd_max_multi, int32_t is_single_end, double min_noise_prob,
                            double prob_precision, int32_t collapse_groups, double * out_noise, uint64_t * out_group_ptr, double * out_group_prob,
                            uint64_t * out_group_id_ptr, uint32_t * out_group_ids);

/* One cluster's PathClusterEstimates for the writers; path p is named "c<cluster_id>_p<p>". */
typedef struct {
    uint32_t cluster_id;
    uint32_t num_paths;
    const uint32_t * length;            /* [num_paths] */
    const double * effective_length;    /* [num_paths] */
    uint64_t n_sets;
    const uint64_t * set_ptr;           /* [n_sets + 1] */
    const uint32_t * set_ids;
    const double * posteriors;          /* [n_sets] */
    uint64_t n_abundances;
    const double * abundances;
    double noise_count;
    double total_count;
} HarnessEstimates;

/* ProbabilityClusterWriter (--write-probs): writes <prefix>.txt.gz with the clusters in the order given. Path lengths:
 * length[c][p], effective_length[c][p] as arrays of pointers per cluster. */
int32_t ref_write_probs(const char * prefix, double prob_precision, uint32_t n, const HarnessCluster * clusters, const uint32_t * const * length,
                        const double * const * effective_length);

/* kind 0: AbundanceEstimatesWriter, 1: HaplotypeAbundanceEstimatesWriter, 2: JointHaplotypeAbundanceEstimatesWriter,
 * 3: JointHaplotypeEstimatesWriter -> <prefix>.txt (with the writers' own header line and, for 0-2, the final "Unknown" line). */
int32_t ref_write_estimates(int32_t kind, const char * prefix, uint32_t ploidy, double min_posterior, double total_transcript_count, uint32_t n,
                            const HarnessEstimates * estimates, uint32_t unaligned_read_count);

#ifdef __cplusplus
}
#endif

#endif
