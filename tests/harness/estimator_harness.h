/* Test harness ABI (test infrastructure, not product).
 *
 * One C++ source, tests/harness/estimator_harness.cpp, is compiled twice:
 *   - against the reference's own classes (oracle/_ref/libharness_ref.so, -DHARNESS_REFERENCE), and
 *   - against this repo's drop-in classes (tests/_build/libharness_b200.so).
 * Both builds expose exactly this ABI, so a parity test is "same call, two libraries".
 * The driver loop in harness_estimate_batch restates rpvg src/main.cpp:811-830, 976-977.
 */
#ifndef RPVG_B200_TESTS_ESTIMATOR_HARNESS_H
#define RPVG_B200_TESTS_ESTIMATOR_HARNESS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One path cluster = the arguments of PathEstimator::estimate (src/path_estimator.hpp:23).
 * Read class r owns entries [entry_ptr[r], entry_ptr[r+1]); entry e is one
 * pair<double, vector<uint32_t>> of ReadPathProbabilities::path_probs
 * (src/read_path_probabilities.hpp:19-44) with ids [entry_path_ptr[e], entry_path_ptr[e+1]). */
typedef struct {
    uint32_t num_paths;
    uint32_t num_reads;
    const uint32_t * read_count;        /* [num_reads] */
    const double * noise_prob;          /* [num_reads] */
    const uint64_t * entry_ptr;         /* [num_reads + 1] */
    const double * entry_prob;          /* [num_entries] */
    const uint64_t * entry_path_ptr;    /* [num_entries + 1] */
    const uint32_t * entry_path_ids;    /* cluster-local path indices */
    const uint32_t * path_group_id;     /* [num_paths] or NULL (-> 0) */
    const uint32_t * path_source_count; /* [num_paths] or NULL (-> 1) */
    const uint64_t * path_source_ptr;   /* [num_paths + 1] or NULL (-> no source ids) */
    const uint32_t * path_source_ids;
} HarnessCluster;

enum {
    HARNESS_MODEL_TRANSCRIPTS = 0,           /* PathAbundanceEstimator         (main.cpp:770) */
    HARNESS_MODEL_STRAINS = 1,               /* MinimumPathAbundanceEstimator  (main.cpp:774) */
    HARNESS_MODEL_HAPLOTYPE_TRANSCRIPTS = 2, /* NestedPathAbundanceEstimator   (main.cpp:778) */
    HARNESS_MODEL_PATH_POSTERIOR = 3,        /* PathPosteriorEstimator */
    HARNESS_MODEL_HAPLOTYPES = 4             /* PathGroupPosteriorEstimator    (main.cpp:766) */
};

typedef struct {
    int32_t model;
    uint32_t group_size;          /* ploidy */
    double min_hap_prob;
    int32_t infer_collapsed;      /* !--ind-hap-inference */
    int32_t use_group_post_gibbs; /* --use-hap-gibbs */
    uint32_t max_em_its;
    double max_rel_em_conv;
    uint32_t num_gibbs_samples;
    uint32_t gibbs_thin_its;
    double prob_precision;
} HarnessConfig;

/* Flattened PathClusterEstimates (src/path_cluster_estimates.hpp:44-110). Owned by the handle. */
typedef struct {
    uint64_t n_sets;
    const uint64_t * set_ptr; /* [n_sets + 1] */
    const uint32_t * set_ids;
    uint64_t n_posteriors;
    const double * posteriors;
    uint64_t n_abundances;
    const double * abundances;
    double noise_count;
    double total_count;
    uint64_t n_gibbs;               /* number of CountSamples */
    const uint64_t * gibbs_path_ptr; /* [n_gibbs + 1] */
    const uint32_t * gibbs_path_ids;
    const uint64_t * gibbs_noise_ptr; /* [n_gibbs + 1] */
    const double * gibbs_noise;
    const uint64_t * gibbs_abund_ptr; /* [n_gibbs + 1] */
    const double * gibbs_abund;
} HarnessResult;

const char * harness_impl_name(void); /* "reference" or "rpvg_b200" */

void * harness_estimator_create(const HarnessConfig * cfg);
void harness_estimator_destroy(void * estimator);

/* estimate() on one cluster with mt19937(seed). Returns a result handle. */
void * harness_estimate(void * estimator, const HarnessCluster * cluster, uint32_t seed);

/* The inference loop of main.cpp:811-830,976-977 over n clusters: largest first (by num_reads),
 * omp parallel for schedule(dynamic,1), mt19937(rng_seed + rank). handles[i] belongs to clusters[i].
 * Returns wall seconds of the loop. num_threads <= 0 -> omp default. */
double harness_estimate_batch(void * estimator, uint32_t n, const HarnessCluster * clusters, uint32_t rng_seed, int32_t num_threads, void ** handles);

#ifndef HARNESS_REFERENCE
/* rpvg_b200 classes only: the same clusters, order and generators through estimateBatch (one call). */
double harness_estimate_batched(void * estimator, uint32_t n, const HarnessCluster * clusters, uint32_t rng_seed, void ** handles);
/* ... through estimateBatchOnDevices: the batch dealt over num_devices GPUs (0: $RPVG_B200_DEVICES). */
double harness_estimate_batched_devices(void * estimator, uint32_t n, const HarnessCluster * clusters, uint32_t rng_seed, int32_t num_devices, void ** handles);
#endif

const HarnessResult * harness_result_view(void * handle);
void harness_result_free(void * handle);

int32_t harness_max_threads(void);
void harness_set_threads(int32_t num_threads);

/* NestedPathAbundanceEstimator::findPathSourceGroups (rpvg:src/path_abundance_estimator.cpp:493-546; host only): path p
 * carries the haplotype ids ids[id_ptr[p] .. id_ptr[p + 1]), inserted in that order. Groups in the order they come out:
 * out_group_ptr [n_groups + 1] into out_group_paths (sized for the sum of all list lengths: at most the number of ids),
 * out_group_counts [n_groups]. Returns n_groups. */
uint32_t harness_source_groups(uint32_t n_paths, const uint64_t * id_ptr, const uint32_t * ids, uint64_t * out_group_ptr, uint32_t * out_group_paths, uint32_t * out_group_counts);

#ifdef HARNESS_REFERENCE
/* ---- kernel-level oracles: the reference's protected members, called one by one ---- */

/* construct{,Partial,Grouped}ProbabilityMatrix (+ addNoiseAndNormalize when add_noise != 0).
 * mode 0: all paths; 1: partial (ids = path_ids, n_ids); 2: grouped (group g = ids[id_ptr[g]..id_ptr[g+1]), n_ids groups).
 * out_matrix: col-major [num_reads x (ncols + (add_noise?1:0))]; out_noise/out_counts: [num_reads]. */
void ref_build_matrix(double prob_precision, const HarnessCluster * cluster, int32_t mode, uint32_t n_ids, const uint64_t * id_ptr, const uint32_t * ids, int32_t add_noise, double * out_matrix, double * out_noise, double * out_counts);

/* sort + quickMergeIdentical scan of the cluster driver (main.cpp:953-973); outputs sized like the inputs. */
uint32_t ref_merge_read_records(double prob_precision, const HarnessCluster * cluster, double * out_noise, uint32_t * out_count, uint64_t * out_entry_ptr, double * out_entry_prob,
                                uint64_t * out_entry_path_ptr, uint32_t * out_entry_path_ids);

/* readCollapseProbabilityMatrix (path_estimator.cpp:219-259). Matrices col-major; returns new row count. */
uint32_t ref_collapse(double prob_precision, uint32_t rows, uint32_t cols, const double * matrix, const double * counts, double * out_matrix, double * out_counts);

/* EMAbundanceEstimator (path_abundance_estimator.cpp:47-114). matrix col-major rows x (n_paths+1). */
void ref_em(uint32_t max_em_its, double max_rel_em_conv, uint32_t rows, uint32_t n_paths, const double * matrix, const double * counts, double * out_abundances, double * out_noise_count, double * out_total_count);

/* EM followed by gibbsReadCountSampler (path_abundance_estimator.cpp:116-212); out_samples [num_samples x n_paths] sample-major. */
void ref_em_gibbs(uint32_t max_em_its, double max_rel_em_conv, uint32_t num_samples, uint32_t thin_its, uint32_t seed, uint32_t rows, uint32_t n_paths, const double * matrix, const double * counts, double * out_abundances, double * out_noise_count, double * out_noise_samples, double * out_abund_samples);

/* calculatePathGroupPosteriors{Full,Bounded} / estimatePathGroupPosteriorsGibbs (path_estimator.cpp:332-589).
 * kind 0 full, 1 bounded, 2 gibbs. matrix col-major rows x n_paths (no noise column). Returns result handle. */
void * ref_group_posteriors(int32_t kind, uint32_t group_size, double min_rel_likelihood, uint32_t seed, uint32_t rows, uint32_t n_paths, const double * matrix, const double * noise, const double * counts, const uint32_t * path_counts);

/* weightedMinimumPathCover (path_abundance_estimator.cpp:297-340). cover col-major rows x n_paths (0/1 bytes). Returns cover size. */
uint32_t ref_path_cover(uint32_t rows, uint32_t n_paths, const uint8_t * cover, const double * counts, const double * weights, uint32_t * out_cover);
#endif

#ifdef __cplusplus
}
#endif

#endif
