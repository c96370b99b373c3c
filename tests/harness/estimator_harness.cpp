// Test harness (test infrastructure, not product). See estimator_harness.h.
// Compiled with -fno-access-control: it fills ReadPathProbabilities' private members directly
// (the only public way to populate them is addPathProbs, which needs alignments + a GBWT).

#include <cstdint>
#include <cstring>
#include <vector>
#include <random>
#include <algorithm>
#include <chrono>
#include <numeric>

#include <omp.h>

#ifdef HARNESS_REFERENCE
#include "path_estimator.hpp"
#include "path_abundance_estimator.hpp"
#include "path_posterior_estimator.hpp"
#include "path_cluster_estimates.hpp"
#include "read_path_probabilities.hpp"
#else
#include "rpvg_b200/host/estimators.hpp"
#endif

#include "estimator_harness.h"

using namespace std;

namespace {

struct ResultHolder {

    vector<uint64_t> set_ptr;
    vector<uint32_t> set_ids;
    vector<double> posteriors;
    vector<double> abundances;
    vector<uint64_t> gibbs_path_ptr;
    vector<uint32_t> gibbs_path_ids;
    vector<uint64_t> gibbs_noise_ptr;
    vector<double> gibbs_noise;
    vector<uint64_t> gibbs_abund_ptr;
    vector<double> gibbs_abund;
    HarnessResult view;
};

ResultHolder * flattenEstimates(const PathClusterEstimates & est) {

    auto * res = new ResultHolder();

    res->set_ptr.push_back(0);
    for (auto & group_set: est.path_group_sets) {
        res->set_ids.insert(res->set_ids.end(), group_set.begin(), group_set.end());
        res->set_ptr.push_back(res->set_ids.size());
    }

    res->posteriors = est.posteriors;
    res->abundances = est.abundances;

    res->gibbs_path_ptr.push_back(0);
    res->gibbs_noise_ptr.push_back(0);
    res->gibbs_abund_ptr.push_back(0);

    for (auto & samples: est.gibbs_read_count_samples) {
        res->gibbs_path_ids.insert(res->gibbs_path_ids.end(), samples.path_ids.begin(), samples.path_ids.end());
        res->gibbs_path_ptr.push_back(res->gibbs_path_ids.size());
        res->gibbs_noise.insert(res->gibbs_noise.end(), samples.noise_samples.begin(), samples.noise_samples.end());
        res->gibbs_noise_ptr.push_back(res->gibbs_noise.size());
        res->gibbs_abund.insert(res->gibbs_abund.end(), samples.abundance_samples.begin(), samples.abundance_samples.end());
        res->gibbs_abund_ptr.push_back(res->gibbs_abund.size());
    }

    auto & v = res->view;
    v.n_sets = est.path_group_sets.size();
    v.set_ptr = res->set_ptr.data();
    v.set_ids = res->set_ids.data();
    v.n_posteriors = res->posteriors.size();
    v.posteriors = res->posteriors.data();
    v.n_abundances = res->abundances.size();
    v.abundances = res->abundances.data();
    v.noise_count = est.noise_count;
    v.total_count = est.total_count;
    v.n_gibbs = est.gibbs_read_count_samples.size();
    v.gibbs_path_ptr = res->gibbs_path_ptr.data();
    v.gibbs_path_ids = res->gibbs_path_ids.data();
    v.gibbs_noise_ptr = res->gibbs_noise_ptr.data();
    v.gibbs_noise = res->gibbs_noise.data();
    v.gibbs_abund_ptr = res->gibbs_abund_ptr.data();
    v.gibbs_abund = res->gibbs_abund.data();

    return res;
}

// What main.cpp:846-907 leaves in the two arguments of estimate(), from flat arrays.
void unflattenCluster(const HarnessCluster & c, double prob_precision, PathClusterEstimates * est, vector<ReadPathProbabilities> * cluster_probs) {

    est->paths.reserve(c.num_paths);

    for (uint32_t p = 0; p < c.num_paths; ++p) {

        est->paths.emplace_back(PathInfo("p" + to_string(p)));
        auto & info = est->paths.back();

        if (c.path_group_id) { info.group_id = c.path_group_id[p]; }
        if (c.path_source_count) { info.source_count = c.path_source_count[p]; }

        if (c.path_source_ptr) {
            for (uint64_t s = c.path_source_ptr[p]; s < c.path_source_ptr[p + 1]; ++s) {
                info.source_ids.emplace(c.path_source_ids[s]);
            }
        }

        info.length = 1000;
        info.effective_length = 800;
    }

    cluster_probs->reserve(c.num_reads);

    for (uint32_t r = 0; r < c.num_reads; ++r) {

        cluster_probs->emplace_back(ReadPathProbabilities(c.read_count[r], prob_precision));
        auto & probs = cluster_probs->back();

        probs.noise_prob = c.noise_prob[r];
        probs.path_probs.reserve(c.entry_ptr[r + 1] - c.entry_ptr[r]);

        for (uint64_t e = c.entry_ptr[r]; e < c.entry_ptr[r + 1]; ++e) {
            probs.path_probs.emplace_back(c.entry_prob[e], vector<uint32_t>(c.entry_path_ids + c.entry_path_ptr[e], c.entry_path_ids + c.entry_path_ptr[e + 1]));
        }
    }
}

struct EstimatorHolder {

    HarnessConfig cfg;
    PathEstimator * estimator;
};

}

extern "C" {

const char * harness_impl_name(void) {

#ifdef HARNESS_REFERENCE
    return "reference";
#else
    return "rpvg_b200";
#endif
}

int32_t harness_max_threads(void) {

    return omp_get_max_threads();
}

// Host threads of the calls that follow (launchers such as torchrun export OMP_NUM_THREADS=1 to their workers).
void harness_set_threads(int32_t num_threads) {

    if (num_threads > 0) { omp_set_num_threads(num_threads); }
}

// findPathSourceGroups of the haplotype-transcripts class (a host routine: no device involved), on paths that carry the
// given haplotype ids, inserted in the given order.
uint32_t harness_source_groups(uint32_t n_paths, const uint64_t * id_ptr, const uint32_t * ids, uint64_t * out_group_ptr, uint32_t * out_group_paths, uint32_t * out_group_counts) {

    vector<PathInfo> paths;
    paths.reserve(n_paths);

    for (uint32_t p = 0; p < n_paths; ++p) {
        paths.emplace_back("path" + to_string(p));
        for (uint64_t k = id_ptr[p]; k < id_ptr[p + 1]; ++k) { paths.back().source_ids.insert(ids[k]); }
        paths.back().source_count = static_cast<uint32_t>(paths.back().source_ids.size());
    }

    NestedPathAbundanceEstimator estimator(2, 0.01, true, false, 100, 0.001, 0, 1, 1e-8);
    const auto groups = estimator.findPathSourceGroups(paths);

    out_group_ptr[0] = 0;
    for (size_t g = 0; g < groups.first.size(); ++g) {
        for (size_t k = 0; k < groups.first[g].size(); ++k) { out_group_paths[out_group_ptr[g] + k] = groups.first[g][k]; }
        out_group_ptr[g + 1] = out_group_ptr[g] + groups.first[g].size();
        out_group_counts[g] = groups.second[g];
    }

    return static_cast<uint32_t>(groups.first.size());
}

// The factory of main.cpp:766-788.
void * harness_estimator_create(const HarnessConfig * cfg) {

    auto * holder = new EstimatorHolder();
    holder->cfg = *cfg;

    switch (cfg->model) {

        case HARNESS_MODEL_TRANSCRIPTS:
            holder->estimator = new PathAbundanceEstimator(cfg->max_em_its, cfg->max_rel_em_conv, cfg->num_gibbs_samples, cfg->gibbs_thin_its, cfg->prob_precision);
            break;
        case HARNESS_MODEL_STRAINS:
            holder->estimator = new MinimumPathAbundanceEstimator(cfg->max_em_its, cfg->max_rel_em_conv, cfg->num_gibbs_samples, cfg->gibbs_thin_its, cfg->prob_precision);
            break;
        case HARNESS_MODEL_HAPLOTYPE_TRANSCRIPTS:
            holder->estimator = new NestedPathAbundanceEstimator(cfg->group_size, cfg->min_hap_prob, cfg->infer_collapsed, cfg->use_group_post_gibbs, cfg->max_em_its, cfg->max_rel_em_conv, cfg->num_gibbs_samples, cfg->gibbs_thin_its, cfg->prob_precision);
            break;
        case HARNESS_MODEL_PATH_POSTERIOR:
            holder->estimator = new PathPosteriorEstimator(cfg->prob_precision);
            break;
        case HARNESS_MODEL_HAPLOTYPES:
            holder->estimator = new PathGroupPosteriorEstimator(cfg->group_size, cfg->use_group_post_gibbs, cfg->prob_precision);
            break;
        default:
            delete holder;
            return nullptr;
    }

    return holder;
}

void harness_estimator_destroy(void * estimator) {

    auto * holder = static_cast<EstimatorHolder *>(estimator);

    if (holder) {
        delete holder->estimator;
        delete holder;
    }
}

void * harness_estimate(void * estimator, const HarnessCluster * cluster, uint32_t seed) {

    auto * holder = static_cast<EstimatorHolder *>(estimator);

    PathClusterEstimates est;
    vector<ReadPathProbabilities> cluster_probs;
    unflattenCluster(*cluster, holder->cfg.prob_precision, &est, &cluster_probs);

    mt19937 mt_rng = mt19937(seed);
    holder->estimator->estimate(&est, cluster_probs, &mt_rng);

    return flattenEstimates(est);
}

double harness_estimate_batch(void * estimator, uint32_t n, const HarnessCluster * clusters, uint32_t rng_seed, int32_t num_threads, void ** handles) {

    auto * holder = static_cast<EstimatorHolder *>(estimator);

    // Inputs are unflattened before the clock starts: in rpvg they come out of
    // ReadPathProbabilities::addPathProbs, which is upstream of the timed path.
    vector<PathClusterEstimates> ests(n);
    vector<vector<ReadPathProbabilities> > probs(n);

    for (uint32_t i = 0; i < n; ++i) {
        unflattenCluster(clusters[i], holder->cfg.prob_precision, &ests[i], &probs[i]);
    }

    vector<pair<uint32_t, uint32_t> > indices;
    indices.reserve(n);

    for (uint32_t i = 0; i < n; ++i) {
        indices.emplace_back(clusters[i].num_reads, i);
    }

    sort(indices.rbegin(), indices.rend());

    if (num_threads > 0) {
        omp_set_num_threads(num_threads);
    }

    auto time_start = chrono::steady_clock::now();

    #pragma omp parallel for schedule(dynamic, 1)
    for (size_t i = 0; i < indices.size(); ++i) {

        auto idx = indices.at(i).second;

        mt19937 mt_rng = mt19937(rng_seed + i);
        holder->estimator->estimate(&ests[idx], probs[idx], &mt_rng);
    }

    auto time_end = chrono::steady_clock::now();

    for (uint32_t i = 0; i < n; ++i) {
        handles[i] = flattenEstimates(ests[i]);
    }

    return chrono::duration<double>(time_end - time_start).count();
}

#ifndef HARNESS_REFERENCE
// The same loop through the batched driver entry of this repository's classes (estimateBatch): same order of clusters,
// same generator per cluster as harness_estimate_batch, one call. Returns wall seconds of the call.
double harness_estimate_batched(void * estimator, uint32_t n, const HarnessCluster * clusters, uint32_t rng_seed, void ** handles) {

    return harness_estimate_batched_devices(estimator, n, clusters, rng_seed, 1, handles);
}

// ... over num_devices GPUs of the box (estimateBatchOnDevices): same inputs, same outputs.
double harness_estimate_batched_devices(void * estimator, uint32_t n, const HarnessCluster * clusters, uint32_t rng_seed, int32_t num_devices, void ** handles) {

    auto * holder = static_cast<EstimatorHolder *>(estimator);

    vector<PathClusterEstimates> ests(n);
    vector<vector<ReadPathProbabilities> > probs(n);

    for (uint32_t i = 0; i < n; ++i) {
        unflattenCluster(clusters[i], holder->cfg.prob_precision, &ests[i], &probs[i]);
    }

    vector<pair<uint32_t, uint32_t> > indices;
    indices.reserve(n);
    for (uint32_t i = 0; i < n; ++i) { indices.emplace_back(clusters[i].num_reads, i); }
    sort(indices.rbegin(), indices.rend());

    vector<mt19937> generators;
    generators.reserve(n);

    vector<PathClusterEstimates *> est_ptrs;
    vector<const vector<ReadPathProbabilities> *> prob_ptrs;
    vector<mt19937 *> rng_ptrs;

    for (size_t i = 0; i < indices.size(); ++i) {
        auto idx = indices.at(i).second;
        generators.emplace_back(rng_seed + i);
        est_ptrs.push_back(&ests[idx]);
        prob_ptrs.push_back(&probs[idx]);
        rng_ptrs.push_back(&generators.back());
    }

    auto time_start = chrono::steady_clock::now();
    if (num_devices == 1) {
        holder->estimator->estimateBatch(est_ptrs, prob_ptrs, rng_ptrs);
    } else {
        holder->estimator->estimateBatchOnDevices(est_ptrs, prob_ptrs, rng_ptrs, num_devices);
    }
    auto time_end = chrono::steady_clock::now();

    for (uint32_t i = 0; i < n; ++i) { handles[i] = flattenEstimates(ests[i]); }

    return chrono::duration<double>(time_end - time_start).count();
}
#endif

const HarnessResult * harness_result_view(void * handle) {

    return &(static_cast<ResultHolder *>(handle)->view);
}

void harness_result_free(void * handle) {

    delete static_cast<ResultHolder *>(handle);
}

#ifdef HARNESS_REFERENCE

void ref_build_matrix(double prob_precision, const HarnessCluster * cluster, int32_t mode, uint32_t n_ids, const uint64_t * id_ptr, const uint32_t * ids, int32_t add_noise, double * out_matrix, double * out_noise, double * out_counts) {

    PathPosteriorEstimator estimator(prob_precision);

    PathClusterEstimates est;
    vector<ReadPathProbabilities> cluster_probs;
    unflattenCluster(*cluster, prob_precision, &est, &cluster_probs);

    Utils::ColMatrixXd read_path_probs;
    Utils::ColVectorXd noise_probs;
    Utils::RowVectorXd read_counts;

    if (mode == 0) {

        estimator.constructProbabilityMatrix(&read_path_probs, &noise_probs, &read_counts, cluster_probs, cluster->num_paths);

    } else if (mode == 1) {

        estimator.constructPartialProbabilityMatrix(&read_path_probs, &noise_probs, &read_counts, cluster_probs, vector<uint32_t>(ids, ids + n_ids), cluster->num_paths);

    } else {

        vector<vector<uint32_t> > groups;

        for (uint32_t g = 0; g < n_ids; ++g) {
            groups.emplace_back(ids + id_ptr[g], ids + id_ptr[g + 1]);
        }

        estimator.constructGroupedProbabilityMatrix(&read_path_probs, &noise_probs, &read_counts, cluster_probs, groups, cluster->num_paths);
    }

    if (add_noise) {
        estimator.addNoiseAndNormalizeProbabilityMatrix(&read_path_probs, noise_probs);
    }

    memcpy(out_matrix, read_path_probs.data(), sizeof(double) * read_path_probs.size());
    memcpy(out_noise, noise_probs.data(), sizeof(double) * noise_probs.size());
    memcpy(out_counts, read_counts.data(), sizeof(double) * read_counts.size());
}

uint32_t ref_collapse(double prob_precision, uint32_t rows, uint32_t cols, const double * matrix, const double * counts, double * out_matrix, double * out_counts) {

    PathPosteriorEstimator estimator(prob_precision);

    Utils::ColMatrixXd read_path_probs = Eigen::Map<const Utils::ColMatrixXd>(matrix, rows, cols);
    Utils::RowVectorXd read_counts = Eigen::Map<const Utils::RowVectorXd>(counts, rows);

    estimator.readCollapseProbabilityMatrix(&read_path_probs, &read_counts);

    memcpy(out_matrix, read_path_probs.data(), sizeof(double) * read_path_probs.size());
    memcpy(out_counts, read_counts.data(), sizeof(double) * read_counts.size());

    return read_path_probs.rows();
}

void ref_em(uint32_t max_em_its, double max_rel_em_conv, uint32_t rows, uint32_t n_paths, const double * matrix, const double * counts, double * out_abundances, double * out_noise_count, double * out_total_count) {

    PathAbundanceEstimator estimator(max_em_its, max_rel_em_conv, 0, 1, 1e-8);

    Utils::ColMatrixXd read_path_probs = Eigen::Map<const Utils::ColMatrixXd>(matrix, rows, n_paths + 1);
    Utils::RowVectorXd read_counts = Eigen::Map<const Utils::RowVectorXd>(counts, rows);

    PathClusterEstimates est;
    est.resetEstimates(n_paths, 1);
    est.total_count = read_counts.sum();

    estimator.EMAbundanceEstimator(&est, read_path_probs, read_counts);

    memcpy(out_abundances, est.abundances.data(), sizeof(double) * n_paths);
    *out_noise_count = est.noise_count;
    *out_total_count = est.total_count;
}

void ref_em_gibbs(uint32_t max_em_its, double max_rel_em_conv, uint32_t num_samples, uint32_t thin_its, uint32_t seed, uint32_t rows, uint32_t n_paths, const double * matrix, const double * counts, double * out_abundances, double * out_noise_count, double * out_noise_samples, double * out_abund_samples) {

    PathAbundanceEstimator estimator(max_em_its, max_rel_em_conv, num_samples, thin_its, 1e-8);

    Utils::ColMatrixXd read_path_probs = Eigen::Map<const Utils::ColMatrixXd>(matrix, rows, n_paths + 1);
    Utils::RowVectorXd read_counts = Eigen::Map<const Utils::RowVectorXd>(counts, rows);

    PathClusterEstimates est;
    est.resetEstimates(n_paths, 1);
    est.total_count = read_counts.sum();

    estimator.EMAbundanceEstimator(&est, read_path_probs, read_counts);

    est.gibbs_read_count_samples.emplace_back(CountSamples());
    est.gibbs_read_count_samples.back().path_ids = vector<uint32_t>(n_paths);
    iota(est.gibbs_read_count_samples.back().path_ids.begin(), est.gibbs_read_count_samples.back().path_ids.end(), 0);

    mt19937 mt_rng = mt19937(seed);
    estimator.gibbsReadCountSampler(&est, read_path_probs, read_counts, 1, &mt_rng, num_samples);

    memcpy(out_abundances, est.abundances.data(), sizeof(double) * n_paths);
    *out_noise_count = est.noise_count;

    memcpy(out_noise_samples, est.gibbs_read_count_samples.back().noise_samples.data(), sizeof(double) * num_samples);
    memcpy(out_abund_samples, est.gibbs_read_count_samples.back().abundance_samples.data(), sizeof(double) * num_samples * n_paths);
}

void * ref_group_posteriors(int32_t kind, uint32_t group_size, double min_rel_likelihood, uint32_t seed, uint32_t rows, uint32_t n_paths, const double * matrix, const double * noise, const double * counts, const uint32_t * path_counts) {

    PathPosteriorEstimator estimator(1e-8);

    Utils::ColMatrixXd read_path_probs = Eigen::Map<const Utils::ColMatrixXd>(matrix, rows, n_paths);
    Utils::ColVectorXd noise_probs = Eigen::Map<const Utils::ColVectorXd>(noise, rows);
    Utils::RowVectorXd read_counts = Eigen::Map<const Utils::RowVectorXd>(counts, rows);

    vector<uint32_t> path_counts_vec(path_counts, path_counts + n_paths);

    PathClusterEstimates est;

    if (kind == 0) {

        estimator.calculatePathGroupPosteriorsFull(&est, read_path_probs, noise_probs, read_counts, path_counts_vec, group_size);

    } else if (kind == 1) {

        estimator.calculatePathGroupPosteriorsBounded(&est, read_path_probs, noise_probs, read_counts, path_counts_vec, group_size, min_rel_likelihood);

    } else {

        mt19937 mt_rng = mt19937(seed);
        estimator.estimatePathGroupPosteriorsGibbs(&est, read_path_probs, noise_probs, read_counts, path_counts_vec, group_size, &mt_rng);
    }

    return flattenEstimates(est);
}

// The cluster driver's sort + merge of the read records (main.cpp:953-973), with the reference's own operator< and
// quickMergeIdentical. Outputs the kept records, flattened: returns their number.
uint32_t ref_merge_read_records(double prob_precision, const HarnessCluster * cluster, double * out_noise, uint32_t * out_count, uint64_t * out_entry_ptr, double * out_entry_prob,
                                uint64_t * out_entry_path_ptr, uint32_t * out_entry_path_ids) {

    PathClusterEstimates unused;
    vector<ReadPathProbabilities> records;
    unflattenCluster(*cluster, prob_precision, &unused, &records);

    sort(records.begin(), records.end());

    if (!records.empty()) {

        uint32_t prev_unique = 0;

        for (size_t i = 1; i < records.size(); ++i) {

            if (!records.at(prev_unique).quickMergeIdentical(records.at(i))) {

                if (prev_unique + 1 < i) { records.at(prev_unique + 1) = records.at(i); }
                prev_unique++;
            }
        }

        records.resize(prev_unique + 1);
    }

    uint64_t n_entries = 0, n_ids = 0;
    out_entry_ptr[0] = 0;
    out_entry_path_ptr[0] = 0;

    for (size_t r = 0; r < records.size(); ++r) {

        out_noise[r] = records[r].noiseProb();
        out_count[r] = records[r].readCount();

        for (auto & group: records[r].pathProbs()) {
            out_entry_prob[n_entries] = group.first;
            for (auto id: group.second) { out_entry_path_ids[n_ids++] = id; }
            out_entry_path_ptr[++n_entries] = n_ids;
        }
        out_entry_ptr[r + 1] = n_entries;
    }

    return static_cast<uint32_t>(records.size());
}

uint32_t ref_path_cover(uint32_t rows, uint32_t n_paths, const uint8_t * cover, const double * counts, const double * weights, uint32_t * out_cover) {

    MinimumPathAbundanceEstimator estimator(1, 1, 0, 1, 1e-8);

    Utils::ColMatrixXb read_path_cover(rows, n_paths);

    for (uint32_t j = 0; j < n_paths; ++j) {
        for (uint32_t i = 0; i < rows; ++i) {
            read_path_cover(i, j) = cover[static_cast<size_t>(j) * rows + i];
        }
    }

    Utils::RowVectorXd read_counts = Eigen::Map<const Utils::RowVectorXd>(counts, rows);
    Utils::RowVectorXd path_weights = Eigen::Map<const Utils::RowVectorXd>(weights, n_paths);

    auto path_cover = estimator.weightedMinimumPathCover(read_path_cover, read_counts, path_weights);
    copy(path_cover.begin(), path_cover.end(), out_cover);

    return path_cover.size();
}

#endif

}
