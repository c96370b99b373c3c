/* rpvg_b200 — C ABI of the B200-native inference hot path of rpvg (everything inside PathEstimator::estimate()).
 *
 * Plain C, plain pointers and sizes, no C++/torch types. Every function returns RPVG_B200_OK (0) or a negative
 * status; rpvg_b200_last_error(ctx) gives the message. Nothing here falls back to the CPU: without a CUDA device
 * rpvg_b200_create fails and every other entry point needs a ctx.
 *
 * Citations "rpvg:<file>:<line>" name the reference code (jonassibbesen/rpvg @ b4d1bdc2) that an entry point replaces.
 *
 * Data layout ("cluster arena"): a batch of mutually independent path clusters (rpvg:src/main.cpp:829 treats them as
 * independent work items), flattened as one CSR over read classes:
 *   cluster c owns read classes [read_offset[c], read_offset[c+1]) and paths (matrix columns) [path_offset[c], path_offset[c+1]);
 *   read class r owns entries [row_ptr[r], row_ptr[r+1]); entry e says "path col[e] (cluster-local index) has probability prob[e]".
 * This is rpvg's vector<ReadPathProbabilities> (rpvg:src/read_path_probabilities.hpp:19-44) with each
 * pair<double, vector<uint32_t>> expanded to one entry per path id and the entries of a read class sorted by path index
 * (strictly ascending: a path occurs at most once per read class), i.e. the non-zeros of the dense row that
 * constructProbabilityMatrix would write (rpvg:src/path_estimator.cpp:55-77).
 */
#ifndef RPVG_B200_H
#define RPVG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RPVG_B200_OK 0
#define RPVG_B200_ERR_INVALID -1  /* bad argument / malformed arena */
#define RPVG_B200_ERR_CUDA -2     /* CUDA runtime error (message has the cudaError string) */
#define RPVG_B200_ERR_NO_DEVICE -3
#define RPVG_B200_ERR_LIMIT -4    /* a size is beyond what an entry point supports */

typedef struct rpvg_b200_ctx rpvg_b200_ctx;     /* per (process, device): stream, scratch pools */
typedef struct rpvg_b200_batch rpvg_b200_batch; /* a cluster arena resident in HBM + its results */

typedef struct {
    uint32_t n_clusters;
    const uint64_t * read_offset; /* [n_clusters + 1] */
    const uint64_t * path_offset; /* [n_clusters + 1] */
    const uint64_t * row_ptr;     /* [n_reads + 1], n_reads = read_offset[n_clusters] */
    const uint32_t * col;         /* [nnz], nnz = row_ptr[n_reads] */
    const double * prob;          /* [nnz]  ReadPathProbabilities::pathProbs() values */
    const double * noise;         /* [n_reads] ReadPathProbabilities::noiseProb() */
    const uint32_t * count;       /* [n_reads] ReadPathProbabilities::readCount() */
} rpvg_b200_clusters;

/* The knobs of PathAbundanceEstimator (rpvg:src/path_abundance_estimator.hpp:22, .cpp:10-14). */
typedef struct {
    uint32_t max_em_its;       /* --max-em-its, default 10000 (rpvg:src/main.cpp:416) */
    double max_rel_em_conv;    /* --max-rel-em-conv, default 0.001 (rpvg:src/main.cpp:417) */
    uint32_t min_em_conv_its;  /* 10   (rpvg:src/path_abundance_estimator.cpp:10) */
    double min_em_abundance;   /* 1e-8 (rpvg:src/path_abundance_estimator.cpp:11) */
} rpvg_b200_em_params;

void rpvg_b200_em_params_default(rpvg_b200_em_params * params);

/* ---- context ---- */

/* device: CUDA ordinal. stream: a cudaStream_t the caller owns (all work of this ctx is enqueued on it), or NULL to
 * let the ctx create its own non-blocking stream. One ctx per host thread that calls in (rpvg calls estimate() from
 * OpenMP threads, rpvg:src/main.cpp:829-830,977); a ctx is not re-entrant. */
int rpvg_b200_create(int device, void * stream, rpvg_b200_ctx ** out_ctx);
/* CUDA devices visible to the process (0 without a usable driver): what the host library shards a batch over. */
int rpvg_b200_device_count(void);
void rpvg_b200_destroy(rpvg_b200_ctx * ctx);
const char * rpvg_b200_last_error(const rpvg_b200_ctx * ctx);
const char * rpvg_b200_version(void);
/* Number of kernel launches this ctx has enqueued so far (for bench.py's gpu_launches). */
uint64_t rpvg_b200_launch_count(const rpvg_b200_ctx * ctx);

/* Page-locked host memory owned by the ctx, for callers that flatten a batch right before handing it over (the batched
 * drivers of rpvg_b200/host): an arena built in it uploads at the full host-link rate instead of going through the
 * driver's staging copy. `slot` < 8 names one grow-only buffer; the pointer stays valid until the next call for the
 * same slot with a larger size, or rpvg_b200_destroy. NULL on failure (rpvg_b200_last_error). No counterpart in the
 * reference (its estimators never leave the host). */
void * rpvg_b200_pinned_scratch(rpvg_b200_ctx * ctx, uint32_t slot, uint64_t bytes);

/* ---- transcripts model: matrix build + EM ----
 * Replaces, for a batch of clusters, the body of PathAbundanceEstimator::estimate
 * (rpvg:src/path_abundance_estimator.cpp:18-33): constructProbabilityMatrix (rpvg:src/path_estimator.cpp:55-77),
 * addNoiseAndNormalizeProbabilityMatrix (:156-166) and EMAbundanceEstimator (rpvg:src/path_abundance_estimator.cpp:47-114).
 * Host pointers in and out; includes the H2D and D2H copies and a stream synchronize.
 *   abundances  [n_paths]    PathClusterEstimates::abundances of every cluster, concatenated (read counts; 0 below min_em_abundance)
 *   noise_count [n_clusters] PathClusterEstimates::noise_count
 *   total_count [n_clusters] PathClusterEstimates::total_count
 *   iterations  [n_clusters] EM iterations executed (may be NULL)
 * Clusters without read classes get zeros (rpvg:src/path_abundance_estimator.cpp:20-22). */
int rpvg_b200_em_abundances(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_em_params * params,
                            double * abundances, double * noise_count, double * total_count, uint32_t * iterations);

/* The same in three steps, for callers that keep the arena resident in HBM (bench.py times step 2 alone):
 * 1. upload: H2D copy of the arena on the ctx stream (returns when the copies have completed, so the host buffers may be
 *    reused; pinned host memory makes the copies run at full PCIe rate).
 * 2. em:     matrix build (row normalisation + noise column + column-major index) and EM; asynchronous.
 * 3. download: D2H of the results + synchronize. */
int rpvg_b200_batch_upload(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, rpvg_b200_batch ** out_batch);
int rpvg_b200_batch_em(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch, const rpvg_b200_em_params * params);
int rpvg_b200_batch_download(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch, double * abundances, double * noise_count, double * total_count, uint32_t * iterations);
void rpvg_b200_batch_free(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch);

/* Sum over clusters of nnz_c (path non-zeros + one noise entry per read class) x iterations_c of the last
 * rpvg_b200_batch_em on this batch, and the algorithmic bytes it moved (12 nnz_path + 20 M + 16 (N+1) per
 * cluster per iteration). Valid after rpvg_b200_batch_download. */
int rpvg_b200_batch_em_work(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch, double * nnz_iterations, double * algorithmic_bytes);

/* Device time (CUDA events on the ctx stream) of the two phases of the last rpvg_b200_batch_em on this batch:
 * matrix build (normalise + column-major index) and the EM kernels. Blocks until that work has finished. */
int rpvg_b200_batch_em_timings(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch, float * build_ms, float * em_ms);

/* ---- haplotype models: log-likelihoods of path-group sets ----
 * The O(#sets x #read classes) core of calculatePathGroupPosteriorsFull (rpvg:src/path_estimator.cpp:332-377),
 * calculatePathGroupPosteriorsBounded (:379-473) and of the conditional distributions of
 * estimatePathGroupPosteriorsGibbs (:527-555): for every set s of group_size paths of its problem (cluster)
 *   loglik[s] = sum_r counts[r] * log(noise[r] + sum_{p in s} matrix[r, p] / group_size).
 * Log frequencies, log(numPermutations), the log-sum-exp and Bounded's sequential pruning are O(#sets) and stay with the
 * caller (rpvg_b200/host does them in the reference's order). Host pointers in and out. A set with a member outside its
 * problem's columns gets NaN. */
typedef struct {
    uint32_t n_problems;
    const uint64_t * row_offset;    /* [n_problems + 1] read classes of problem p */
    const uint64_t * path_offset;   /* [n_problems + 1] paths (matrix columns) of problem p */
    const uint64_t * matrix_offset; /* [n_problems + 1] into matrix; problem p owns rows_p x paths_p doubles, column-major */
    const uint64_t * set_offset;    /* [n_problems + 1] sets of problem p */
    const double * matrix;
    const double * noise;           /* [rows] */
    const double * counts;          /* [rows] read counts (as doubles: collapsed classes carry sums) */
    uint32_t group_size;            /* ploidy k, 1..8 */
    const uint32_t * sets;          /* [n_sets * group_size] problem-local path indices */
} rpvg_b200_loglik_problems;

int rpvg_b200_group_loglik(rpvg_b200_ctx * ctx, const rpvg_b200_loglik_problems * problems, double * loglik);

/* Micro-benchmark: fp64 log() evaluations per second this GPU sustains (a log, an add and a fused multiply-add per
 * evaluation, no memory traffic; a few milliseconds). The bound of every kernel that takes a log per term, as the
 * reference's formulation does (no counterpart in the reference; MEASURED_PEAKS.json has no such figure). */
int rpvg_b200_log_peak(rpvg_b200_ctx * ctx, double * logs_per_second);

/* Micro-benchmark: fp64 fused multiply-adds per second (independent chains, no memory traffic). The pair kernel of
 * rpvg_b200_grouped_bounded_pairs spends two fp64 instructions per term (the sum noise + a / 2 + b / 2 and the product it
 * goes into) instead of a log: half of this figure is its roofline (no counterpart in the reference). */
int rpvg_b200_fp64_peak(rpvg_b200_ctx * ctx, double * fma_per_second);

/* ---- read-class collapse ----
 * rowSortProbabilityMatrix + readCollapseProbabilityMatrix (rpvg:src/path_estimator.cpp:13-31,197-259) for a batch of
 * problems: rows of a normalised probability matrix (sparse: entries with problem-local column indices ascending, an
 * absent column is 0; the noise column is the separate per-row value, compared last) are sorted with the reference's
 * tolerance comparator and merged greedily against the running representative (all columns within prob_precision).
 * Exact: the partition of rows into classes, the class counts and the class order.
 *   row_class   [n_rows]       class (index in output order, within its problem) of every input row; may be NULL
 *   n_classes   [n_problems]
 *   class_rep   [>= sum n_classes] input row whose values the class keeps (classes of problem p start at the sum of
 *               n_classes over earlier problems); may be NULL. Capacity n_rows always suffices.
 *   class_count [>= sum n_classes] summed read counts; may be NULL */
typedef struct {
    uint32_t n_problems;
    const uint64_t * row_offset; /* [n_problems + 1] */
    const uint64_t * row_ptr;    /* [n_rows + 1] */
    const uint32_t * col;
    const double * val;
    const double * noise;        /* [n_rows] */
    const double * count;        /* [n_rows] */
} rpvg_b200_collapse_problems;

int rpvg_b200_collapse(rpvg_b200_ctx * ctx, const rpvg_b200_collapse_problems * problems, double prob_precision, uint32_t * row_class, uint32_t * n_classes,
                       uint32_t * class_rep, double * class_count);

/* ---- EM on path subsets: the many small EMs of the haplotype-transcripts model, the restricted EM of the strains model ----
 * For every subset s of the paths of cluster subset_cluster[s]: constructPartialProbabilityMatrix
 * (rpvg:src/path_estimator.cpp:79-113), addNoiseAndNormalizeProbabilityMatrix (:156-166), readCollapseProbabilityMatrix
 * (:219-259) and EMAbundanceEstimator (rpvg:src/path_abundance_estimator.cpp:47-114) — the body of the loop over sampled
 * haplotype combinations at rpvg:src/path_abundance_estimator.cpp:625-671 (and of :260-269 for the minimum path cover) —
 * for all subsets of all clusters in one batched device pipeline. Host pointers in and out.
 *   abundances  [sum |subset|]  per subset path, in subset order (read counts; 0 below min_em_abundance)
 *   noise_count, total_count [n_subsets]; iterations [n_subsets] (may be NULL)
 *   n_classes   [n_subsets] read classes left after the collapse (may be NULL) */
typedef struct {
    uint32_t n_subsets;
    const uint32_t * subset_cluster; /* [n_subsets] cluster of the arena each subset restricts */
    const uint64_t * subset_offset;  /* [n_subsets + 1] into subset_paths */
    const uint32_t * subset_paths;   /* cluster-local path indices, strictly ascending within a subset (duplicates removed) */
} rpvg_b200_subsets;

int rpvg_b200_subset_em(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_subsets * subsets, const rpvg_b200_em_params * params,
                        double prob_precision, double * abundances, double * noise_count, double * total_count, uint32_t * iterations, uint32_t * n_classes);

/* ---- haplotype-transcripts model: the grouped, normalised, collapsed matrix the haplotype posteriors are computed on ----
 * Per cluster: constructGroupedProbabilityMatrix (rpvg:src/path_estimator.cpp:115-154: column g = sum over the paths of
 * group g), addNoiseAndNormalizeProbabilityMatrix (:156-166), readCollapseProbabilityMatrix (:219-259), noise column
 * split off again (rpvg:src/path_abundance_estimator.cpp:440-446). With one group per path of a transcript it is the
 * partial-matrix variant of inferAbundancesIndependentGroups (:377-383).
 *   n_classes [n_clusters]   read classes left per cluster
 *   matrix    capacity sum_c M_c x G_c; cluster c's block starts at sum_{c' < c} M_c' x G_c' and holds
 *             n_classes[c] x G_c doubles, column-major (leading dimension n_classes[c])
 *   noise, counts capacity n_reads; cluster c's values start at read_offset[c] (first n_classes[c] entries) */
typedef struct {
    const uint64_t * group_offset;   /* [n_clusters + 1] groups of cluster c */
    const uint64_t * group_path_ptr; /* [n_groups + 1] */
    const uint32_t * group_paths;    /* cluster-local path indices */
} rpvg_b200_path_groups;

int rpvg_b200_grouped_collapse(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_path_groups * groups, double prob_precision,
                               uint32_t * n_classes, double * matrix, double * noise, double * counts);

/* The same matrices kept on the device, for callers that go on to rpvg_b200_group_loglik with them (the batched driver
 * of the haplotype-transcripts class): rpvg_b200_grouped_create = rpvg_b200_grouped_collapse without the copy back,
 * rpvg_b200_grouped_loglik = rpvg_b200_group_loglik on those matrices (problem c = cluster c, paths = its groups;
 * set_offset [n_clusters + 1], sets problem-local group indices). Free with rpvg_b200_grouped_free. */
typedef struct rpvg_b200_grouped rpvg_b200_grouped;

int rpvg_b200_grouped_create(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_path_groups * groups, double prob_precision,
                             uint32_t * n_classes, rpvg_b200_grouped ** grouped);
int rpvg_b200_grouped_loglik(rpvg_b200_ctx * ctx, const rpvg_b200_grouped * grouped, uint32_t group_size, const uint64_t * set_offset, const uint32_t * sets,
                             double * loglik);
/* The raw matrices of the haplotypes model (constructProbabilityMatrix, rpvg:src/path_estimator.cpp:55-77: the
 * probabilities as they are, noise apart, neither normalised nor collapsed) built on the device from the arena and kept
 * there: problem c = cluster c, its columns = its paths. Every cluster needs at least one read class. */
int rpvg_b200_grouped_raw(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, rpvg_b200_grouped ** grouped);

/* Dense matrices the caller built itself (the layout of rpvg_b200_loglik_problems; its `sets` is not used), put on the
 * device once for several rounds of rpvg_b200_grouped_loglik: the batched driver of the haplotype posterior classes. */
int rpvg_b200_grouped_upload(rpvg_b200_ctx * ctx, const rpvg_b200_loglik_problems * problems, rpvg_b200_grouped ** grouped);
/* The pair walk of calculatePathGroupPosteriorsBounded (rpvg:src/path_estimator.cpp:424-450) on the resident matrices:
 * for each cluster all pairs (i, j >= i) of order[] (its groups in the caller's order, at group_offset[c]) get
 * ll = rpvg_b200_grouped_loglik's sum + log_freq[a] + log_freq[b] + log(numPermutations); a pair is a candidate unless
 * ll - (largest ll of the pairs before it) < min_log_diff. Returned: per cluster the candidates in pair order —
 * cand_pair = index of the pair in that enumeration, cand_loglik = its sum over the read classes (without the
 * frequency terms) — in memory owned by the handle (valid until the next call on it or its release). Pass
 * min_log_diff a little below log(min_rel_likelihood) and replay the reference's walk on the candidates: pairs left
 * out fail the exact test as well and never hold the running maximum. */
int rpvg_b200_grouped_bounded_pairs(rpvg_b200_ctx * ctx, rpvg_b200_grouped * grouped, const uint64_t * group_offset, const uint32_t * order, const double * log_freq,
                                    double min_log_diff, const uint64_t ** cand_offset, const uint32_t ** cand_pair, const double ** cand_loglik);
/* Device time (CUDA events around the pair kernels) and number of terms (pairs x read classes: a log each in the
 * reference's formulation) summed over the rpvg_b200_grouped_bounded_pairs calls made on this handle: what bench.py's
 * roofline of the pair kernel is computed from. */
int rpvg_b200_grouped_pair_stats(const rpvg_b200_grouped * grouped, double * kernel_ms, double * log_evaluations);
void rpvg_b200_grouped_free(rpvg_b200_ctx * ctx, rpvg_b200_grouped * grouped);

/* ---- strains model: greedy weighted minimum path cover ----
 * The cover construction of MinimumPathAbundanceEstimator::estimate (rpvg:src/path_abundance_estimator.cpp:229-251:
 * cover = the non-zero pattern, read counts zeroed for noise-only read classes, weight_j = -sum_i log(p_ij) count_i) and
 * weightedMinimumPathCover (:297-340) for every cluster of the arena. in_cover [n_paths]: 1 for the paths picked.
 * (A cluster with one path gets {0}; a cluster whose read classes are all noise-only gets the empty cover.) */
int rpvg_b200_path_cover(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, uint8_t * in_cover);

/* MinimumPathAbundanceEstimator::weightedMinimumPathCover itself (rpvg:src/path_abundance_estimator.cpp:297-340; the public
 * method the reference's own test calls, rpvg:src/tests/path_abundance_estimator_test.cpp:8-28) on a cover the caller made:
 * read class r covers the paths col[row_ptr[r] .. row_ptr[r+1]) (the true entries of row r of the boolean matrix),
 * read_counts[n_reads] and path_weights[n_paths] as the reference takes them. in_cover[n_paths]: 1 for the picked paths
 * (ascending index = the reference's sorted result). One cluster per call. */
int rpvg_b200_weighted_path_cover(rpvg_b200_ctx * ctx, uint32_t n_reads, uint32_t n_paths, const uint64_t * row_ptr, const uint32_t * col,
                                  const double * read_counts, const double * path_weights, uint8_t * in_cover);

/* ---- Gibbs read-count sampler ----
 * gibbsReadCountSampler (rpvg:src/path_abundance_estimator.cpp:116-212), started from the EM result of each cluster
 * (or path subset): num_samples[c] samples, one every thin_its iterations. Random numbers are counter-based
 * (Philox4x32-10 keyed by seeds[c]); the reference's sequential mt19937 stream is not reproduced, results agree in
 * distribution and are bit-reproducible for equal seeds.
 *   abundance_samples: cluster c's block starts at sum_{c' < c} num_samples[c'] * N_c' and is sample-major
 *                      (CountSamples::abundance_samples); noise_samples: block at sum_{c' < c} num_samples[c']. */
typedef struct {
    const uint32_t * num_samples; /* [n_clusters] or [n_subsets]; 0 = none for that cluster */
    uint32_t thin_its;            /* --gibbs-thin-its, default 25 (rpvg:src/main.cpp:418) */
    const uint64_t * seeds;       /* [n_clusters] or [n_subsets] */
    double * abundance_samples;   /* out */
    double * noise_samples;       /* out */
} rpvg_b200_gibbs;

/* rpvg_b200_em_abundances / rpvg_b200_subset_em followed by the sampler (gibbs may be NULL: then they are the same calls). */
int rpvg_b200_em_abundances_gibbs(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_em_params * params, const rpvg_b200_gibbs * gibbs,
                                  double * abundances, double * noise_count, double * total_count, uint32_t * iterations);
int rpvg_b200_subset_em_gibbs(rpvg_b200_ctx * ctx, const rpvg_b200_clusters * clusters, const rpvg_b200_subsets * subsets, const rpvg_b200_em_params * params,
                              double prob_precision, const rpvg_b200_gibbs * gibbs, double * abundances, double * noise_count, double * total_count,
                              uint32_t * iterations, uint32_t * n_classes);
/* The sampler alone, on a resident batch whose EM has run. */
int rpvg_b200_batch_gibbs(rpvg_b200_ctx * ctx, rpvg_b200_batch * batch, const uint32_t * num_samples, uint32_t thin_its, const uint64_t * seeds,
                          double * abundance_samples, double * noise_samples);

/* ---- sort + merge of the read records of every cluster, before they reach the estimators ----
 * What rpvg:src/main.cpp:953-973 does per cluster with std::sort (operator< of ReadPathProbabilities,
 * rpvg:src/read_path_probabilities.cpp:283-322) and quickMergeIdentical (:223-250), for all clusters of a batch.
 * A record = noise probability, read count and its probability groups (probability + ascending path ids), as
 * ReadPathProbabilities::pathProbs() holds them.
 *   n_merged[c]                                : records cluster c keeps
 *   merged_rep[read_offset[c] + k], k < n_merged: cluster-local index of the k-th kept record, in the reference's order
 *   merged_count[...same...]                   : its read count after the merges (uint32 arithmetic like the reference)
 *   record_class[r]                            : k of the kept record that input record r went into
 * Exact: the partition, the order and the counts. Records the comparator calls equal (all values within 100 eps, same
 * count) are in arbitrary order in the reference (unstable std::sort) and in input order here. */
typedef struct rpvg_b200_read_records {
    uint32_t n_clusters;
    const uint64_t * read_offset;   /* [n_clusters + 1] first record of each cluster */
    const uint64_t * group_ptr;     /* [n_records + 1] first probability group of each record */
    const double * group_prob;      /* [n_groups] */
    const uint64_t * id_ptr;        /* [n_groups + 1] first path id of each group */
    const uint32_t * ids;           /* path ids */
    const double * noise;           /* [n_records] */
    const uint32_t * count;         /* [n_records] */
} rpvg_b200_read_records;

int rpvg_b200_merge_read_records(rpvg_b200_ctx * ctx, const rpvg_b200_read_records * records, double prob_precision, uint32_t * n_merged, uint32_t * record_class,
                                 uint32_t * merged_rep, uint32_t * merged_count);

/* rpvg_b200_grouped_raw (the raw dense matrices of the haplotypes model, constructProbabilityMatrix,
 * rpvg:src/path_estimator.cpp:55-77) from the read records as ReadPathProbabilities::pathProbs() holds them — a
 * probability and the paths that share it — instead of the expanded arena: a third of the bytes cross the host link.
 * path_offset [n_clusters + 1]: paths of every cluster. */
int rpvg_b200_grouped_raw_records(rpvg_b200_ctx * ctx, const rpvg_b200_read_records * records, const uint64_t * path_offset, rpvg_b200_grouped ** grouped);

/* ---- per-read path probabilities: the step right before the estimators' boundary ----
 * ReadPathProbabilities::addPathProbs (rpvg:src/read_path_probabilities.cpp:74-221) with calcAlignPathLogProbs (:39-67)
 * for all reads of a batch of clusters. Read r (of the cluster whose read_offset range holds it) owns the alignment paths
 * align_ptr[r] .. align_ptr[r+1]; the LAST of them is the noise alignment path (AlignmentPath with an empty search state:
 * score_sum <= 0, no path ids), every other one lists the paths it was found on (ids[id_ptr[a] .. id_ptr[a+1]), indices
 * into the cluster's paths, i.e. after rpvg's clustered_path_index lookup; any order, repeats across alignment paths
 * allowed). frag_log_prob[a] = FragmentLengthDist::logProb(frag_length) of alignment path a (rpvg:src/fragment_length_dist.cpp:385;
 * the distribution stays with the caller), NULL for single-end reads. path_group != NULL switches collapse_groups on:
 * path_group[p] = index of the path's group among the groups of its cluster (rpvg's group_name_index), source_count[p] its
 * weight (NULL = 1); the output ids are then group indices.
 * Output, for read r: noise[r] = ReadPathProbabilities::noiseProb(), and pathProbs() as groups group_ptr[r] .. group_ptr[r+1]
 * — group g has the probability group_prob[g] and the ascending ids group_ids[group_id_ptr[g] .. group_id_ptr[g+1]) — i.e.
 * the layout rpvg_b200_read_records takes (read counts are the caller's: 1 per read before the merge). Capacities:
 * noise [n_reads], group_ptr [n_reads + 1], group_prob and group_ids [number of input ids], group_id_ptr [that + 1].
 * Same arithmetic as the reference except that exp/log/log1p are the device's: values agree to ~1e-14 relative. A read
 * with mapping quality 0 keeps noise 1 and no groups; one whose noise alignment path has score 0 gets its noise (~1) and
 * no groups; malformed input (index out of range, a read whose paths all have effective length 0, which rpvg asserts
 * against) is RPVG_B200_ERR_INVALID. Host pointers in and out. */
typedef struct {
    uint32_t n_clusters;
    const uint64_t * read_offset;      /* [n_clusters + 1] reads of cluster c */
    const uint64_t * path_offset;      /* [n_clusters + 1] paths of cluster c */
    const double * effective_length;   /* [n_paths] PathInfo::effective_length */
    const uint32_t * path_group;       /* [n_paths] or NULL */
    const uint32_t * source_count;     /* [n_paths] or NULL */
    const uint8_t * min_mapq;          /* [n_reads] AlignmentPath::min_mapq (equal over the alignment paths of a read) */
    const uint64_t * align_ptr;        /* [n_reads + 1] */
    const int32_t * score_sum;         /* [n_align] AlignmentPath::score_sum */
    const uint32_t * align_length;     /* [n_align] AlignmentPath::align_length */
    const double * frag_log_prob;      /* [n_align] or NULL */
    const uint64_t * id_ptr;           /* [n_align + 1] */
    const uint32_t * ids;              /* cluster-local path indices */
} rpvg_b200_alignments;

int rpvg_b200_add_path_probs(rpvg_b200_ctx * ctx, const rpvg_b200_alignments * alignments, double min_noise_prob, double prob_precision, double * noise,
                             uint64_t * group_ptr, double * group_prob, uint64_t * group_id_ptr, uint32_t * group_ids);

/* ---- transcripts per million (the reducer behind the abundance table) ----
 * total_transcript_count = sum over all paths with effective_length > 0 of abundance / effective_length
 * (rpvg:src/main.cpp:1029-1057); tpm[i] = abundance[i] / effective_length[i] / total_transcript_count * 10^6, 0 for a
 * path without effective length (AbundanceEstimatesWriter::addEstimates, rpvg:src/threaded_output_writer.cpp:305-320).
 * Arrays are over all paths of all clusters of a run (host memory); total_transcript_count may be NULL. As in the
 * reference, a run without any transcript count divides by zero (NaN). */
int rpvg_b200_tpm(rpvg_b200_ctx * ctx, uint64_t n_paths, const double * abundances, const double * effective_lengths, double * tpm,
                  double * total_transcript_count);

/* ---- per-path haplotype probabilities and read counts (the reducer behind the haplotype abundance table) ----
 * HaplotypeAbundanceEstimatesWriter::addEstimates, rpvg:src/threaded_output_writer.cpp:355-391, for all clusters of a run:
 * the estimates of cluster c are the path sets set_offset[c] .. set_offset[c+1]; set i has the members
 * members[member_offset[i] .. member_offset[i+1]) (cluster-local path indices, at most `ploidy`, repeats adjacent),
 * the posterior posteriors[i] and one abundance per member (abundances is parallel to members).
 *   read_counts[p]     = sum of the abundances of the members equal to p
 *   haplotype_probs[p] = sum of the posteriors of the sets p is a member of (a member equal to its predecessor in the set
 *                        adds nothing)
 * both over the paths of all clusters (path_offset[c] + local index), added in the order of the sets: bit-identical to
 * the reference's loop. Host memory in and out. */
int rpvg_b200_haplotype_marginals(rpvg_b200_ctx * ctx, uint32_t n_clusters, const uint64_t * path_offset, const uint64_t * set_offset, const uint64_t * member_offset,
                                  const uint32_t * members, const double * posteriors, const double * abundances, double * haplotype_probs, double * read_counts);

/* ---- read counts and TPM per member of every haplotype combination (the reducer behind the joint haplotype abundance table) ----
 * JointHaplotypeAbundanceEstimatesWriter::addEstimates, rpvg:src/threaded_output_writer.cpp:456-525, for all clusters of a
 * run; sets, members and posteriors as in rpvg_b200_haplotype_marginals. Only sets with posterior >= min_posterior are
 * written (written[i] = 1), and — like the reference's iterator, :473-502 — the abundances of a cluster are consumed by the
 * members of its WRITTEN sets only: abundances[member_offset[set_offset[c]] + k] belongs to the k-th member of cluster c's
 * written sets (with every set written this is the k-th member, the layout of rpvg_b200_haplotype_marginals).
 *   read_counts[e], tpm[e] : per member e of a written set (0 for the others): that abundance, and
 *                            (abundance / effective_lengths[path], 0 without effective length) / total_transcript_count * 10^6
 *   noise_counts[ploidy]   : each slot = sum over the clusters, in order, of noise_count[c] / ploidy (:517-520)
 * total_transcript_count is rpvg_b200_tpm's. Bit-identical to the reference's arithmetic. Host memory in and out. */
int rpvg_b200_joint_haplotype_estimates(rpvg_b200_ctx * ctx, uint32_t n_clusters, const uint64_t * path_offset, const uint64_t * set_offset,
                                        const uint64_t * member_offset, const uint32_t * members, const double * posteriors, const double * abundances,
                                        const double * effective_lengths, const double * noise_count, uint32_t ploidy, double min_posterior,
                                        double total_transcript_count, uint8_t * written, double * read_counts, double * tpm, double * noise_counts);

/* ---- EM of one cluster for callers that must not allocate ----
 * No counterpart in the reference (rpvg runs on the host and allocates freely); it exists for callers of this library
 * that cannot call cudaMalloc while they run: code under a fixed device-memory budget sized at start-up, real-time
 * loops, or hosts living inside another framework's allocator. Unlike every other entry point it allocates nothing
 * (no cudaMalloc/cudaFree/cudaMallocAsync, no allocating library call): inputs and outputs are DEVICE pointers owned by
 * the caller and the kernel's work area is a fixed 4 MiB array inside the library's device image (held by every GPU that
 * loads the library until exit).
 *   limit   : 8 * (nnz + 2 * (n_paths + 1)) <= 4 MiB, else RPVG_B200_ERR_LIMIT (message in error_message) and no launch.
 *   inputs  : one cluster in the arena layout above: d_row_ptr[n_reads + 1] (entry offsets, any base), d_col/d_prob
 *             indexed by those offsets, d_noise[n_reads], d_count[n_reads]; nnz = d_row_ptr[n_reads] - d_row_ptr[0].
 *   outputs : d_abundances[n_paths], d_noise_count[1], d_total_count[1], d_iterations[1]. The kernel is enqueued on
 *             `stream` (a cudaStream_t, NULL = default stream); the call does not synchronize and leaves `device` the
 *             calling thread's current device. Same semantics as rpvg_b200_em_abundances for one cluster, equal to it
 *             up to rounding.
 *   checks  : made on the device, like the regular entry point's: nnz must equal d_row_ptr[n_reads] - d_row_ptr[0] (the
 *             host sizes the work area from it), offsets non-decreasing, path indices < n_paths and strictly ascending
 *             within a read class, probabilities finite and >= 0. On failure d_iterations[0] = 0xffffffff and nothing
 *             else is written (neither the outputs nor the work area).
 *   sharing : all calls on a device use the same work area and find the previous call's contents in it (the kernel clears
 *             what it accumulates into); calls whose kernels could overlap on the GPU race — serialise them per device.
 *   set-up  : the first call on a device resolves the work area's device address (the runtime may load the library's
 *             device code on that GPU at this point); make it before any stream capture. */
int rpvg_b200_em_cluster_noalloc(int device, void * stream, uint32_t n_reads, uint32_t n_paths, uint64_t nnz, const uint64_t * d_row_ptr,
                                 const uint32_t * d_col, const double * d_prob, const double * d_noise, const uint32_t * d_count,
                                 const rpvg_b200_em_params * params, double * d_abundances, double * d_noise_count, double * d_total_count,
                                 uint32_t * d_iterations, char * error_message, size_t error_message_len);

#ifdef __cplusplus
}
#endif

#endif
