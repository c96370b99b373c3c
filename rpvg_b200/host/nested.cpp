// NestedPathAbundanceEstimator (haplotype-transcripts model) on top of the C ABI.
//
// What runs on the GPU: the grouped / partial matrix + normalisation + read-class collapse
// (rpvg_b200_grouped_collapse), the log-likelihood of every haplotype combination (rpvg_b200_group_loglik, via the base
// class) and ALL sub-EMs of a cluster — one per sampled haplotype combination — in one batched call
// (rpvg_b200_subset_em). What stays here is the combinatorics whose order the output depends on: the reference keeps
// haplotype groups, path subsets and the per-transcript estimates in sparsepp hash maps and emits them in iteration
// order (rpvg:src/path_abundance_estimator.cpp:508-542, 625, 738), so the same containers, keys, hash and insertion
// sequence are used.

#include "estimators.hpp"

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <iostream>

#include "rpvg_b200.h"

namespace {

[[noreturn]] void fatalNested(rpvg_b200_ctx * ctx, const char * what) {

    cerr << "ERROR: rpvg_b200 " << what << ": " << rpvg_b200_last_error(ctx) << endl;
    abort();
}

// same flattening as in estimators.cpp (one cluster as a one-cluster arena)
struct FlatCluster {

    vector<uint64_t> read_offset, path_offset, row_ptr;
    vector<uint32_t> col, count;
    vector<double> prob, noise;

    FlatCluster(const vector<ReadPathProbabilities> & cluster_probs, const uint32_t num_paths) {

        read_offset = {0, cluster_probs.size()};
        path_offset = {0, num_paths};
        row_ptr = {0};

        vector<pair<uint32_t, double> > row;

        for (auto & read_probs: cluster_probs) {

            row.clear();
            for (auto & path_probs: read_probs.pathProbs()) {
                for (auto & path: path_probs.second) { row.emplace_back(path, path_probs.first); }
            }
            sort(row.begin(), row.end());

            for (auto & entry: row) {
                col.push_back(entry.first);
                prob.push_back(entry.second);
            }

            row_ptr.push_back(col.size());
            noise.push_back(read_probs.noiseProb());
            count.push_back(read_probs.readCount());
        }
    }

    rpvg_b200_clusters view() const {
        return rpvg_b200_clusters{1, read_offset.data(), path_offset.data(), row_ptr.data(), col.data(), prob.data(), noise.data(), count.data()};
    }
};

}

NestedPathAbundanceEstimator::NestedPathAbundanceEstimator(const uint32_t group_size_in, const double min_hap_prob_in, const bool infer_collapsed_in, const bool use_group_post_gibbs_in, const uint32_t max_em_its, const double max_rel_em_conv, const uint32_t num_gibbs_samples, const uint32_t gibbs_thin_its, const double prob_precision) : PathAbundanceEstimator(max_em_its, max_rel_em_conv, num_gibbs_samples, gibbs_thin_its, prob_precision), group_size(group_size_in), min_hap_prob(min_hap_prob_in), infer_collapsed(infer_collapsed_in), use_group_post_gibbs(use_group_post_gibbs_in) {}

void NestedPathAbundanceEstimator::estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) {

    if (infer_collapsed) {
        inferAbundancesCollapsedGroups(path_cluster_estimates, cluster_probs, mt_rng);
    } else {
        inferAbundancesIndependentGroups(path_cluster_estimates, cluster_probs, mt_rng);
    }
}

// Posteriors of the size-k multisets of `path_groups` (rpvg:src/path_abundance_estimator.cpp:377-409 and 440-464).
void NestedPathAbundanceEstimator::groupPosteriors(PathClusterEstimates * group_estimates, const vector<ReadPathProbabilities> & cluster_probs, const uint32_t num_paths, const vector<vector<uint32_t> > & path_groups, const vector<uint32_t> & group_counts, mt19937 * mt_rng) const {

    const FlatCluster flat(cluster_probs, num_paths);
    rpvg_b200_clusters clusters = flat.view();

    vector<uint64_t> group_offset = {0, path_groups.size()};
    vector<uint64_t> group_path_ptr = {0};
    vector<uint32_t> group_paths;

    for (auto & group: path_groups) {
        group_paths.insert(group_paths.end(), group.begin(), group.end());
        group_path_ptr.push_back(group_paths.size());
    }

    rpvg_b200_path_groups groups = {group_offset.data(), group_path_ptr.data(), group_paths.data()};

    const size_t rows = cluster_probs.size();
    const size_t num_groups = path_groups.size();

    uint32_t n_classes = 0;
    vector<double> matrix(rows * num_groups), noise(rows), counts(rows);

    auto * ctx = threadContext();

    if (rpvg_b200_grouped_collapse(ctx, &clusters, &groups, prob_precision, &n_classes, matrix.data(), noise.data(), counts.data()) != RPVG_B200_OK) {
        fatalNested(ctx, "rpvg_b200_grouped_collapse");
    }

    matrix.resize(static_cast<size_t>(n_classes) * num_groups);
    noise.resize(n_classes);
    counts.resize(n_classes);

    if (use_group_post_gibbs) {

        estimatePathGroupPosteriorsGibbs(group_estimates, matrix, noise, counts, num_groups, group_counts, group_size, mt_rng);

    } else if (group_size == 2) {

        calculatePathGroupPosteriorsBounded(group_estimates, matrix, noise, counts, num_groups, group_counts, group_size, min_hap_prob);

    } else {

        calculatePathGroupPosteriorsFull(group_estimates, matrix, noise, counts, num_groups, group_counts, group_size);
    }
}

// rpvg:src/path_abundance_estimator.cpp:473-491 (group order = first appearance of the transcript id)
vector<vector<uint32_t> > NestedPathAbundanceEstimator::findPathGroups(const vector<PathInfo> & paths) const {

    vector<vector<uint32_t> > path_groups;
    rpvg_b200_host::hash_map<uint32_t, uint32_t> path_group_indexes;

    for (size_t i = 0; i < paths.size(); ++i) {

        auto inserted = path_group_indexes.emplace(paths.at(i).group_id, path_group_indexes.size());
        if (inserted.second) { path_groups.emplace_back(vector<uint32_t>()); }

        path_groups.at(inserted.first->second).emplace_back(i);
    }

    return path_groups;
}

// rpvg:src/path_abundance_estimator.cpp:493-546: haplotype id -> its paths; haplotypes with the same path list form one
// group whose count is its multiplicity. Groups come out in the hash map's iteration order (of their first haplotype).
pair<vector<vector<uint32_t> >, vector<uint32_t> > NestedPathAbundanceEstimator::findPathSourceGroups(const vector<PathInfo> & paths) const {

    // The reference's map from haplotype id to the paths that carry it (rpvg:src/path_abundance_estimator.cpp:495-504:
    // emplace + push_back per (path, id)) is only ever walked in its iteration order, which depends on the sequence of
    // insertions alone, i.e. on the order in which the ids are first met — not on the mapped type. So the lists are
    // gathered in one flat array by counting (no hashing and no allocation per (path, id) pair: a 1000GP-scale batch has
    // millions of them), and a map from id to id receives every id once, in first-met order: same keys in the same
    // sequence, same order of iteration.
    // (the id sets are walked once, in their own order; everything after that runs over a flat copy)
    size_t n_pairs = 0;
    for (auto & path: paths) { n_pairs += path.source_ids.size(); }

    vector<uint32_t> pair_id;
    vector<uint32_t> pair_path;
    pair_id.reserve(n_pairs);
    pair_path.reserve(n_pairs);

    uint32_t max_id = 0;
    for (size_t i = 0; i < paths.size(); ++i) {
        for (auto & id: paths[i].source_ids) {
            max_id = max(max_id, id);
            pair_id.push_back(id);
            pair_path.push_back(static_cast<uint32_t>(i));
        }
    }

    pair<vector<vector<uint32_t> >, vector<uint32_t> > path_source_groups;

    if (n_pairs > 0 && static_cast<size_t>(max_id) <= 64 * n_pairs + 1024) {

        vector<uint32_t> begin(static_cast<size_t>(max_id) + 2, 0);
        vector<uint32_t> first_met;

        for (auto id: pair_id) {
            if (begin[id + 1]++ == 0) { first_met.push_back(id); }
        }

        for (size_t id = 0; id + 1 < begin.size(); ++id) { begin[id + 1] += begin[id]; }

        vector<uint32_t> flat(n_pairs);
        {
            vector<uint32_t> cursor(begin.begin(), begin.end() - 1);
            for (size_t k = 0; k < n_pairs; ++k) { flat[cursor[pair_id[k]]++] = pair_path[k]; }
        }

        rpvg_b200_host::hash_map<uint32_t, uint32_t> id_order;
        for (auto id: first_met) { id_order.emplace(id, id); }

        // One pass instead of the reference's comparison of every list with every later one: a group is emitted when its
        // path list is first met in the map's iteration order, every later haplotype with that list adds one to its count.
        // (lists are compared through a 64-bit hash first; its map is only looked up, never walked)
        std::unordered_map<uint64_t, vector<uint32_t> > groups_of_hash;

        for (auto it = id_order.begin(); it != id_order.end(); ++it) {

            const uint32_t * list = flat.data() + begin[it->first];
            const size_t len = begin[it->first + 1] - begin[it->first];

            uint64_t h = 1469598103934665603ull ^ len;
            for (size_t k = 0; k < len; ++k) { h = (h ^ list[k]) * 1099511628211ull; }

            auto & candidates = groups_of_hash[h];
            bool found = false;

            for (auto g: candidates) {
                const auto & have = path_source_groups.first[g];
                if (have.size() == len && equal(have.begin(), have.end(), list)) {
                    path_source_groups.second[g]++;
                    found = true;
                    break;
                }
            }

            if (!found) {
                candidates.push_back(static_cast<uint32_t>(path_source_groups.first.size()));
                path_source_groups.first.emplace_back(list, list + len);
                path_source_groups.second.emplace_back(1);
            }
        }

        return path_source_groups;
    }

    // (ids too sparse for a table: the reference's own sequence of operations)
    rpvg_b200_host::hash_map<uint32_t, vector<uint32_t> > source_id_paths;

    for (size_t i = 0; i < paths.size(); ++i) {
        for (auto & id: paths.at(i).source_ids) {
            auto inserted = source_id_paths.emplace(id, vector<uint32_t>());
            inserted.first->second.emplace_back(i);
        }
    }

    rpvg_b200_host::hash_map<vector<uint32_t>, uint32_t> group_of_list;

    for (auto it = source_id_paths.begin(); it != source_id_paths.end(); ++it) {

        if (it->second.empty()) { continue; }

        auto inserted = group_of_list.emplace(it->second, path_source_groups.first.size());

        if (inserted.second) {
            path_source_groups.first.emplace_back(it->second);
            path_source_groups.second.emplace_back(1);
        } else {
            path_source_groups.second.at(inserted.first->second)++;
        }
    }

    return path_source_groups;
}

// selectPathSubsetIndices (rpvg:src/path_abundance_estimator.cpp:569-606): combinations at or above min_hap_prob, expanded
// to sorted path lists, their posteriors renormalised
NestedPathAbundanceEstimator::SubsetSamples NestedPathAbundanceEstimator::collapsedSubsetSamples(const PathClusterEstimates & group_estimates, const vector<vector<uint32_t> > & path_groups) const {

    SubsetSamples path_subset_samples;
    double sum_posterior = 0;

    for (size_t i = 0; i < group_estimates.posteriors.size(); ++i) {

        if (group_estimates.posteriors.at(i) >= min_hap_prob) {

            vector<uint32_t> path_subset;
            for (auto & group: group_estimates.path_group_sets.at(i)) {
                for (auto & path: path_groups.at(group)) { path_subset.emplace_back(path); }
            }

            sort(path_subset.begin(), path_subset.end());

            auto inserted = path_subset_samples.emplace(path_subset, 0);
            inserted.first->second += group_estimates.posteriors.at(i);

            sum_posterior += group_estimates.posteriors.at(i);
        }
    }

    for (auto & subset_sample: path_subset_samples) { subset_sample.second /= sum_posterior; }

    return path_subset_samples;
}

// rpvg:src/path_abundance_estimator.cpp:428-471
void NestedPathAbundanceEstimator::inferAbundancesCollapsedGroups(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) const {

    path_cluster_estimates->resetEstimates(0, 0);

    if (cluster_probs.empty()) { return; }

    auto path_source_groups = findPathSourceGroups(path_cluster_estimates->paths);

    PathClusterEstimates group_estimates;
    groupPosteriors(&group_estimates, cluster_probs, path_cluster_estimates->paths.size(), path_source_groups.first, path_source_groups.second, mt_rng);

    const SubsetSamples path_subset_samples = collapsedSubsetSamples(group_estimates, path_source_groups.first);

    inferPathSubsetAbundance(path_cluster_estimates, cluster_probs, mt_rng, path_subset_samples);
}

// rpvg:src/path_abundance_estimator.cpp:356-426
void NestedPathAbundanceEstimator::inferAbundancesIndependentGroups(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) const {

    path_cluster_estimates->resetEstimates(0, 0);

    if (cluster_probs.empty()) { return; }

    auto path_groups = findPathGroups(path_cluster_estimates->paths);

    vector<vector<uint32_t> > path_subset_samples(floor(1 / min_hap_prob));

    for (auto & group: path_groups) {

        // one column per path of the transcript: the partial matrix of the reference
        vector<vector<uint32_t> > singletons;
        vector<uint32_t> group_path_counts;

        for (auto & path: group) {
            singletons.emplace_back(vector<uint32_t>({path}));
            group_path_counts.emplace_back(path_cluster_estimates->paths.at(path).source_count);
        }

        PathClusterEstimates group_estimates;
        groupPosteriors(&group_estimates, cluster_probs, path_cluster_estimates->paths.size(), singletons, group_path_counts, mt_rng);

        // sampleGroupPathIndices (:548-567): one draw per joint sample, host RNG
        discrete_distribution<uint32_t> path_group_set_sampler(group_estimates.posteriors.begin(), group_estimates.posteriors.end());

        for (auto & path_subset_sample: path_subset_samples) {

            vector<uint32_t> path_group_set = group_estimates.path_group_sets.at(path_group_set_sampler(*mt_rng));
            sort(path_group_set.begin(), path_group_set.end());

            for (auto & path_group: path_group_set) { path_subset_sample.emplace_back(group.at(path_group)); }
        }
    }

    SubsetSamples clustered_path_subset_samples;

    for (auto & path_subset: path_subset_samples) {

        sort(path_subset.begin(), path_subset.end());

        auto inserted = clustered_path_subset_samples.emplace(path_subset, 0);
        inserted.first->second += 1 / static_cast<double>(path_subset_samples.size());
    }

    inferPathSubsetAbundance(path_cluster_estimates, cluster_probs, mt_rng, clustered_path_subset_samples);
}

// rpvg:src/path_abundance_estimator.cpp:608-750, with the sub-EMs of all subsets batched into one device call.
vector<NestedPathAbundanceEstimator::SubsetJob> NestedPathAbundanceEstimator::planPathSubsets(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng, const SubsetSamples & path_subset_samples) const {

    for (auto & cluster_prob: cluster_probs) { path_cluster_estimates->total_count += cluster_prob.readCount(); }

    // pass 1, in the map's iteration order: the subsets that take part, duplicates (homozygous paths) removed.
    // :675-683 the sample budget is split over the subsets by conditional binomials, host generator, subset order
    uint32_t subset_gibbs_samples = num_gibbs_samples;
    double subset_gibbs_prob = 1;

    vector<SubsetJob> jobs;

    for (auto & path_subset: path_subset_samples) {

        if (path_subset.second < min_hap_prob) { continue; }

        SubsetJob job;
        job.paths = path_subset.first;
        job.prob = path_subset.second;

        for (auto & path: path_subset.first) {
            if (job.collapsed.empty() || job.collapsed.back() != path) {
                job.collapsed.emplace_back(path);
                job.multiplicity.emplace_back(1);
            } else {
                job.multiplicity.back()++;
            }
        }

        job.gibbs_samples = 0;
        job.gibbs_seed = 0;

        if (subset_gibbs_samples > 0) {

            binomial_distribution<uint32_t> sample_count_sampler(subset_gibbs_samples, min(1.0, path_subset.second / subset_gibbs_prob));
            job.gibbs_samples = sample_count_sampler(*mt_rng);

            subset_gibbs_samples -= job.gibbs_samples;
            subset_gibbs_prob -= path_subset.second;
        }

        jobs.emplace_back(std::move(job));
    }

    // keys of the device sampler: after the allotment, in subset order
    for (auto & job: jobs) {
        if (job.gibbs_samples > 0) { job.gibbs_seed = (static_cast<uint64_t>((*mt_rng)()) << 32) | (*mt_rng)(); }
    }

    return jobs;
}

void NestedPathAbundanceEstimator::finishPathSubsets(PathClusterEstimates * path_cluster_estimates, const vector<SubsetJob> & jobs, const double * abundances, const double * noise_counts, const double * gibbs_abundances, const double * gibbs_noise) const {

    // :685-696 one CountSamples per subset with samples, in subset order
    size_t value_pos = 0, sample_pos = 0;

    for (auto & job: jobs) {

        if (job.gibbs_samples > 0) {

            CountSamples samples;
            samples.path_ids = job.collapsed;
            samples.noise_samples.assign(gibbs_noise + sample_pos, gibbs_noise + sample_pos + job.gibbs_samples);
            samples.abundance_samples.assign(gibbs_abundances + value_pos, gibbs_abundances + value_pos + static_cast<size_t>(job.gibbs_samples) * job.collapsed.size());

            path_cluster_estimates->gibbs_read_count_samples.emplace_back(std::move(samples));
        }

        sample_pos += job.gibbs_samples;
        value_pos += static_cast<size_t>(job.gibbs_samples) * job.collapsed.size();
    }

    // pass 2, same order: weight by the subset's probability and file under the transcript's path tuple
    rpvg_b200_host::hash_map<vector<uint32_t>, pair<double, vector<double> > > path_group_estimates;
    double sum_hap_prob = 0;
    size_t subset_pos = 0;

    for (size_t s = 0; s < jobs.size(); ++s) {

        const auto & job = jobs.at(s);
        const double * subset_abundances = abundances + subset_pos;
        subset_pos += job.collapsed.size();

        sum_hap_prob += job.prob;
        path_cluster_estimates->noise_count += noise_counts[s] * job.prob;

        rpvg_b200_host::hash_map<uint32_t, vector<uint32_t> > subset_path_group_index;

        for (auto & path: job.paths) {
            auto inserted = subset_path_group_index.emplace(path_cluster_estimates->paths.at(path).group_id, vector<uint32_t>());
            inserted.first->second.emplace_back(path);
        }

        for (auto & path_group: subset_path_group_index) {

            assert(path_group.second.size() <= group_size);

            auto inserted = path_group_estimates.emplace(path_group.second, pair<uint32_t, vector<double> >(0, vector<double>(path_group.second.size(), 0)));
            inserted.first->second.first += job.prob;

            for (size_t i = 0; i < path_group.second.size(); ++i) {

                const size_t idx = lower_bound(job.collapsed.begin(), job.collapsed.end(), path_group.second.at(i)) - job.collapsed.begin();
                inserted.first->second.second.at(i) += (subset_abundances[idx] * job.prob / job.multiplicity.at(idx));
            }
        }
    }

    path_cluster_estimates->path_group_sets.reserve(path_group_estimates.size());
    path_cluster_estimates->posteriors.reserve(path_group_estimates.size());
    path_cluster_estimates->abundances.reserve(path_group_estimates.size() * group_size);

    for (auto & estimates: path_group_estimates) {

        path_cluster_estimates->path_group_sets.emplace_back(estimates.first);
        path_cluster_estimates->posteriors.emplace_back(estimates.second.first);
        path_cluster_estimates->abundances.insert(path_cluster_estimates->abundances.end(), estimates.second.second.begin(), estimates.second.second.end());
    }

    path_cluster_estimates->noise_count += (1 - sum_hap_prob) * path_cluster_estimates->total_count;
}

void NestedPathAbundanceEstimator::inferPathSubsetAbundance(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng, const SubsetSamples & path_subset_samples) const {

    const vector<SubsetJob> jobs = planPathSubsets(path_cluster_estimates, cluster_probs, mt_rng, path_subset_samples);

    // all sub-EMs of the cluster in one call
    vector<double> abundances, gibbs_abundances, gibbs_noise;
    vector<double> noise_counts(jobs.size(), 0), total_counts(jobs.size(), 0);

    if (!jobs.empty()) {

        const FlatCluster flat(cluster_probs, path_cluster_estimates->paths.size());
        rpvg_b200_clusters clusters = flat.view();

        vector<uint32_t> subset_cluster(jobs.size(), 0);
        vector<uint64_t> subset_offset = {0};
        vector<uint32_t> subset_paths;
        vector<uint32_t> gibbs_counts;
        vector<uint64_t> gibbs_seeds;
        size_t total_samples = 0, total_values = 0;

        for (auto & job: jobs) {
            subset_paths.insert(subset_paths.end(), job.collapsed.begin(), job.collapsed.end());
            subset_offset.push_back(subset_paths.size());
            gibbs_counts.push_back(job.gibbs_samples);
            gibbs_seeds.push_back(job.gibbs_seed);
            total_samples += job.gibbs_samples;
            total_values += static_cast<size_t>(job.gibbs_samples) * job.collapsed.size();
        }

        rpvg_b200_subsets subsets = {static_cast<uint32_t>(jobs.size()), subset_cluster.data(), subset_offset.data(), subset_paths.data()};

        rpvg_b200_em_params params;
        rpvg_b200_em_params_default(&params);
        params.max_em_its = max_em_its;
        params.max_rel_em_conv = max_rel_em_conv;

        abundances.resize(subset_paths.size());
        gibbs_abundances.resize(total_values);
        gibbs_noise.resize(total_samples);

        rpvg_b200_gibbs gibbs = {gibbs_counts.data(), gibbs_thin_its, gibbs_seeds.data(), gibbs_abundances.data(), gibbs_noise.data()};

        auto * ctx = threadContext();

        if (rpvg_b200_subset_em_gibbs(ctx, &clusters, &subsets, &params, prob_precision, total_samples > 0 ? &gibbs : nullptr, abundances.data(), noise_counts.data(),
                                      total_counts.data(), nullptr, nullptr) != RPVG_B200_OK) {
            fatalNested(ctx, "rpvg_b200_subset_em");
        }
    }

    finishPathSubsets(path_cluster_estimates, jobs, abundances.data(), noise_counts.data(), gibbs_abundances.data(), gibbs_noise.data());
}
