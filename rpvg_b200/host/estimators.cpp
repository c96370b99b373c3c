// Host side of the drop-in estimators: flattening, the O(#sets) combinatorics in the reference's order, calls into the
// C ABI. No numerical work over the data happens here; without a GPU threadContext() aborts (rpvg's own error style:
// asserts are live, rpvg:CMakeLists.txt:180).

#include "estimators.hpp"

#include <condition_variable>
#include <mutex>

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <limits>
#include <numeric>

#include "rpvg_b200.h"

namespace {

// rpvg:src/utils.hpp:300-302
inline double add_log(double log_x, double log_y) {

    return log_x > log_y ? log_x + log1p(exp(log_y - log_x)) : log_y + log1p(exp(log_x - log_y));
}

// rpvg:src/utils.hpp:95-117 — kept as it is (not the multinomial coefficient for k >= 4)
inline uint32_t numPermutations(vector<uint32_t> values) {

    assert(!values.empty());
    if (values.size() == 1) { return 1; }

    sort(values.begin(), values.end());

    uint32_t num_unique_values = 1;
    for (size_t i = 1; i < values.size(); ++i) {
        if (values.at(i - 1) != values.at(i)) { num_unique_values++; }
    }

    return (tgamma(values.size() + 1) / tgamma(values.size() - num_unique_values + 2));
}

// rpvg:src/path_estimator.cpp:315-330
vector<double> calcPathLogFrequences(const vector<uint32_t> & path_counts) {

    uint32_t count_sum = accumulate(path_counts.begin(), path_counts.end(), 0);
    assert(count_sum > 0);

    vector<double> path_log_freqs;
    path_log_freqs.reserve(path_counts.size());

    for (auto & count: path_counts) {
        assert(count > 0);
        path_log_freqs.emplace_back(log(count / static_cast<double>(count_sum)));
    }

    return path_log_freqs;
}

void generateGroupsRecursive(vector<vector<uint32_t> > * sets, const uint32_t num_components, const uint32_t group_size, vector<uint32_t> cur_group) {

    if (cur_group.size() < group_size) {

        const uint32_t start_idx = cur_group.empty() ? 0 : cur_group.back();

        for (uint32_t i = start_idx; i < num_components; ++i) {
            vector<uint32_t> new_group = cur_group;
            new_group.push_back(i);
            generateGroupsRecursive(sets, num_components, group_size, new_group);
        }

    } else {

        sets->emplace_back(cur_group);
    }
}

[[noreturn]] void fatal(rpvg_b200_ctx * ctx, const char * what) {

    cerr << "ERROR: rpvg_b200 " << what << ": " << rpvg_b200_last_error(ctx) << endl;
    abort();
}

// One cluster as a one-cluster arena: every pair<double, vector<uint32_t>> expanded, entries of a read class sorted by path.
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

// One cluster's dense matrix put on the device once (rpvg_b200_grouped_upload), for every log-likelihood request that follows.
struct ResidentCluster {

    rpvg_b200_ctx * ctx;
    rpvg_b200_grouped * handle = nullptr;
    uint32_t num_paths;

    ResidentCluster(rpvg_b200_ctx * ctx_in, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths_in) : ctx(ctx_in), num_paths(num_paths_in) {

        const uint64_t rows = noise.size();
        const uint64_t row_offset[2] = {0, rows};
        const uint64_t path_offset[2] = {0, num_paths};
        const uint64_t matrix_offset[2] = {0, rows * num_paths};
        const uint64_t set_offset[2] = {0, 0};

        rpvg_b200_loglik_problems problems;
        problems.n_problems = 1;
        problems.row_offset = row_offset;
        problems.path_offset = path_offset;
        problems.matrix_offset = matrix_offset;
        problems.set_offset = set_offset;
        problems.matrix = matrix.data();
        problems.noise = noise.data();
        problems.counts = counts.data();
        problems.group_size = 1;
        problems.sets = nullptr;

        if (rpvg_b200_grouped_upload(ctx, &problems, &handle) != RPVG_B200_OK) { fatal(ctx, "rpvg_b200_grouped_upload"); }
    }

    ~ResidentCluster() { rpvg_b200_grouped_free(ctx, handle); }

    ResidentCluster(const ResidentCluster &) = delete;
    ResidentCluster & operator=(const ResidentCluster &) = delete;

    vector<double> logLikelihoods(const uint32_t group_size, const vector<uint32_t> & flat_sets) const {

        const uint64_t set_offset[2] = {0, flat_sets.size() / group_size};
        vector<double> loglik(set_offset[1], 0);
        if (set_offset[1] > 0 && rpvg_b200_grouped_loglik(ctx, handle, group_size, set_offset, flat_sets.data(), loglik.data()) != RPVG_B200_OK) { fatal(ctx, "rpvg_b200_grouped_loglik"); }
        return loglik;
    }
};

// sum_r count_r log(noise_r + sum_p col_p / k) for each set, on the device
vector<double> deviceLogLikelihoods(rpvg_b200_ctx * ctx, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts,
                                    const uint32_t num_paths, const uint32_t group_size, const vector<uint32_t> & flat_sets) {

    const uint64_t rows = noise.size();
    const uint64_t n_sets = flat_sets.size() / group_size;

    const uint64_t row_offset[2] = {0, rows};
    const uint64_t path_offset[2] = {0, num_paths};
    const uint64_t matrix_offset[2] = {0, rows * num_paths};
    const uint64_t set_offset[2] = {0, n_sets};

    rpvg_b200_loglik_problems problems;
    problems.n_problems = 1;
    problems.row_offset = row_offset;
    problems.path_offset = path_offset;
    problems.matrix_offset = matrix_offset;
    problems.set_offset = set_offset;
    problems.matrix = matrix.data();
    problems.noise = noise.data();
    problems.counts = counts.data();
    problems.group_size = group_size;
    problems.sets = flat_sets.data();

    vector<double> loglik(n_sets, 0);
    if (n_sets > 0 && rpvg_b200_group_loglik(ctx, &problems, loglik.data()) != RPVG_B200_OK) { fatal(ctx, "rpvg_b200_group_loglik"); }

    return loglik;
}

}

#ifndef RPVG_B200_WITH_RPVG_HEADERS

void PathClusterEstimates::resetEstimates(uint32_t num_components, const uint32_t group_size) {

    path_group_sets.clear();

    posteriors.clear();
    abundances.clear();

    noise_count = 0;
    total_count = 0;

    gibbs_read_count_samples.clear();

    if (group_size > 0) {

        generateGroupsRecursive(&path_group_sets, num_components, group_size, vector<uint32_t>());

        posteriors = vector<double>(path_group_sets.size(), 0);
        abundances = vector<double>(path_group_sets.size() * group_size, 0);
    }
}

#endif

// Callers of estimate() that found the device busy. One mutex: a request is two pointers, the queue is touched for
// nanoseconds around device calls of hundreds of microseconds. Up to kLeaders batches are under way at a time, each on
// its own context: the host stages of one overlap the device time of another, and a batch that holds one slowly
// converging cluster does not keep every caller waiting (RPVG_B200_COMBINE_SLOTS, default 3).
struct PathEstimator::Combiner {

    struct Request {
        PathClusterEstimates * estimates;
        const vector<ReadPathProbabilities> * cluster_probs;
        mt19937 * rng;
        bool taken = false;   // a leader has it in its batch
        bool done = false;
    };

    static constexpr size_t kMaxSlots = 8;

    mutex lock;
    condition_variable wake;
    vector<Request *> waiting;
    const bool enabled;
    size_t n_slots = 3;
    bool slot_busy[kMaxSlots] = {false, false, false, false, false, false, false, false};
    rpvg_b200_ctx * slot_ctx[kMaxSlots] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

    Combiner() : enabled(!(getenv("RPVG_B200_COMBINE") && string(getenv("RPVG_B200_COMBINE")) == "0")) {
        if (const char * v = getenv("RPVG_B200_COMBINE_SLOTS")) { n_slots = min<size_t>(kMaxSlots, max<size_t>(1, strtoul(v, nullptr, 10))); }
    }

    ~Combiner() {
        for (auto ctx: slot_ctx) { if (ctx) { rpvg_b200_destroy(ctx); } }
    }

    // (under the lock) index of a free slot, or n_slots
    size_t freeSlot() const {
        size_t k = 0;
        while (k < n_slots && slot_busy[k]) { ++k; }
        return k;
    }
};

PathEstimator::PathEstimator(const double prob_precision_in) : combiner(new Combiner()), prob_precision(prob_precision_in) {}

PathEstimator::~PathEstimator() { delete combiner; }

void PathEstimator::estimate(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) {

    if (!combiner->enabled || !combinesCallers()) {
        estimateOne(path_cluster_estimates, cluster_probs, mt_rng);
        return;
    }

    Combiner::Request mine;
    mine.estimates = path_cluster_estimates;
    mine.cluster_probs = &cluster_probs;
    mine.rng = mt_rng;

    unique_lock<mutex> guard(combiner->lock);
    combiner->waiting.push_back(&mine);

    for (;;) {

        // until somebody has done this request, or a slot is free and this thread does the waiting ones itself
        combiner->wake.wait(guard, [&]() { return mine.done || (!mine.taken && combiner->freeSlot() < combiner->n_slots); });
        if (mine.done) { return; }

        const size_t slot = combiner->freeSlot();
        combiner->slot_busy[slot] = true;

        vector<Combiner::Request *> batch;
        batch.swap(combiner->waiting);   // this thread's request is among them: queued before the wait, taken by nobody
        for (auto request: batch) { request->taken = true; }

        guard.unlock();

        // One context per slot serves all combined calls (only one thread is ever in a slot): the calling threads need
        // none of their own.
        if (!combiner->slot_ctx[slot]) {
            const int status = rpvg_b200_create(threadDevice(), nullptr, &combiner->slot_ctx[slot]);
            if (status != RPVG_B200_OK) {
                cerr << "ERROR: rpvg_b200_create(device " << threadDevice() << ") failed with status " << status << " (a CUDA device is required; there is no CPU fallback)" << endl;
                abort();
            }
        }

        rpvg_b200_ctx * previous = contextOverride();
        contextOverride() = combiner->slot_ctx[slot];

        if (batch.size() == 1) {

            estimateOne(batch[0]->estimates, *batch[0]->cluster_probs, batch[0]->rng);

        } else {

            vector<PathClusterEstimates *> est;
            vector<const vector<ReadPathProbabilities> *> probs;
            vector<mt19937 *> generators;

            for (auto request: batch) {
                est.push_back(request->estimates);
                probs.push_back(request->cluster_probs);
                generators.push_back(request->rng);
            }

            estimateBatch(est, probs, generators);
        }

        contextOverride() = previous;

        guard.lock();

        for (auto request: batch) { request->done = true; }
        combiner->slot_busy[slot] = false;
        combiner->wake.notify_all();
    }
}

rpvg_b200_ctx *& PathEstimator::contextOverride() {

    thread_local rpvg_b200_ctx * ctx = nullptr;
    return ctx;
}

namespace {

// The contexts of one host thread, one per device it has worked on; released when the thread ends.
struct ThreadContexts {

    static constexpr int kMaxDevices = 64;
    rpvg_b200_ctx * ctx[kMaxDevices] = {};

    ~ThreadContexts() {
        for (auto & c: ctx) {
            if (c) { rpvg_b200_destroy(c); }
            c = nullptr;
        }
    }
};

}

int & PathEstimator::threadDevice() {

    thread_local int device = -1;

    if (device < 0) {
        const char * env = getenv("RPVG_B200_DEVICE");
        device = env ? atoi(env) : 0;
    }

    return device;
}

rpvg_b200_ctx * PathEstimator::threadContext() {

    if (contextOverride()) { return contextOverride(); }

    thread_local ThreadContexts contexts;

    const int device = threadDevice();

    if (device < 0 || device >= ThreadContexts::kMaxDevices) {
        cerr << "ERROR: rpvg_b200: device ordinal " << device << " out of range" << endl;
        abort();
    }

    if (!contexts.ctx[device]) {

        const int status = rpvg_b200_create(device, nullptr, &contexts.ctx[device]);

        if (status != RPVG_B200_OK) {
            cerr << "ERROR: rpvg_b200_create(device " << device << ") failed with status " << status << " (a CUDA device is required; there is no CPU fallback)" << endl;
            abort();
        }
    }

    return contexts.ctx[device];
}

// entries x expected EM iterations; the iterations grow roughly with the square root of the cluster's size (config 2:
// ~30 for a handful of entries, ~1000 at 1e4, 3000-5000 at 1e5), the stopping rule caps them
double PathEstimator::clusterCost(const PathClusterEstimates & estimates, const vector<ReadPathProbabilities> & cluster_probs) const {

    double weight = cluster_probs.size();
    for (auto & read_probs: cluster_probs) {
        for (auto & path_probs: read_probs.pathProbs()) { weight += path_probs.second.size(); }
    }

    return weight * min(5000.0, max(20.0, 12.0 * sqrt(weight)));
}

void PathEstimator::denseMatrix(vector<double> * matrix, vector<double> * noise, vector<double> * counts, const vector<ReadPathProbabilities> & cluster_probs, const uint32_t num_paths) {

    const size_t rows = cluster_probs.size();

    matrix->assign(rows * num_paths, 0);
    noise->resize(rows);
    counts->resize(rows);

    for (size_t i = 0; i < rows; ++i) {

        for (auto & path_probs: cluster_probs.at(i).pathProbs()) {
            for (auto & path: path_probs.second) {
                assert(path < num_paths);
                (*matrix)[static_cast<size_t>(path) * rows + i] = path_probs.first;
            }
        }

        (*noise)[i] = cluster_probs.at(i).noiseProb();
        (*counts)[i] = cluster_probs.at(i).readCount();
    }
}

void PathEstimator::planPathGroupPosteriors(GroupPosteriorPlan * plan, PathClusterEstimates * path_cluster_estimates, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size, const bool bounded, const double min_rel_likelihood) const {

    assert(group_size > 0);
    assert(!bounded || group_size == 2);

    plan->num_paths = num_paths;
    plan->group_size = group_size;
    plan->bounded = bounded;
    plan->min_log_likelihood_diff = bounded ? log(min_rel_likelihood) : 0;
    plan->path_log_freqs = calcPathLogFrequences(path_counts);
    plan->stage = 0;
    plan->marginal_posteriors.clear();

    // Full: every size-k multiset. Bounded (rpvg:src/path_estimator.cpp:398): the marginals first, i.e. Full with k = 1.
    PathClusterEstimates * first = bounded ? &plan->marginal_estimates : path_cluster_estimates;

    if (bounded) { path_cluster_estimates->resetEstimates(0, 0); }
    first->resetEstimates(num_paths, bounded ? 1 : group_size);

    plan->set_size = bounded ? 1 : group_size;
    plan->flat_sets.clear();
    plan->flat_sets.reserve(first->path_group_sets.size() * plan->set_size);
    for (auto & group_set: first->path_group_sets) { plan->flat_sets.insert(plan->flat_sets.end(), group_set.begin(), group_set.end()); }
}

bool PathEstimator::advancePathGroupPosteriors(GroupPosteriorPlan * plan, PathClusterEstimates * path_cluster_estimates, const double * loglik) const {

    if (plan->stage == 0) {

        // rpvg:src/path_estimator.cpp:332-377 from the log-likelihoods on
        PathClusterEstimates * estimates = plan->bounded ? &plan->marginal_estimates : path_cluster_estimates;

        double sum_log_posterior = numeric_limits<double>::lowest();

        for (uint32_t i = 0; i < estimates->path_group_sets.size(); ++i) {

            estimates->posteriors.at(i) = loglik[i];

            for (auto & path_idx: estimates->path_group_sets.at(i)) { estimates->posteriors.at(i) += plan->path_log_freqs.at(path_idx); }

            estimates->posteriors.at(i) += log(numPermutations(estimates->path_group_sets.at(i)));
            sum_log_posterior = add_log(sum_log_posterior, estimates->posteriors.at(i));
        }

        for (size_t i = 0; i < estimates->posteriors.size(); ++i) {
            estimates->posteriors.at(i) = exp(estimates->posteriors.at(i) - sum_log_posterior);
        }

        if (!plan->bounded) {
            plan->stage = 2;
            return false;
        }

        // rpvg:src/path_estimator.cpp:400-412: paths by marginal posterior, descending (ties: index descending)
        plan->marginal_posteriors.clear();
        plan->marginal_posteriors.reserve(estimates->posteriors.size());

        for (size_t i = 0; i < estimates->posteriors.size(); ++i) {
            plan->marginal_posteriors.emplace_back(estimates->posteriors.at(i), estimates->path_group_sets.at(i).front());
        }

        sort(plan->marginal_posteriors.rbegin(), plan->marginal_posteriors.rend());

        // Every pair (i, j >= i) in the reference's enumeration order is evaluated on the device; the reference's pruning
        // (per-first-path upper bound, running maximum) is then replayed over the finished values. A pair the bound would
        // have skipped has a log-likelihood below the bound, so the replay drops it as well: the list comes out the same.
        plan->set_size = 2;
        plan->flat_sets.clear();

        if (plan->device_pairs) {   // the caller enumerates the pairs on the device from marginal_posteriors
            plan->stage = 1;
            return true;
        }

        plan->flat_sets.reserve(static_cast<size_t>(plan->num_paths) * (plan->num_paths + 1));

        for (uint32_t i = 0; i < plan->marginal_posteriors.size(); ++i) {
            for (uint32_t j = i; j < plan->marginal_posteriors.size(); ++j) {
                plan->flat_sets.push_back(plan->marginal_posteriors.at(i).second);
                plan->flat_sets.push_back(plan->marginal_posteriors.at(j).second);
            }
        }

        plan->stage = 1;
        return true;
    }

    assert(plan->stage == 1);

    vector<double> log_likelihoods;
    double max_log_likelihood = numeric_limits<double>::lowest();

    size_t pair_idx = 0;

    for (uint32_t i = 0; i < plan->marginal_posteriors.size(); ++i) {

        const uint32_t first_path_idx = plan->marginal_posteriors.at(i).second;

        for (uint32_t j = i; j < plan->marginal_posteriors.size(); ++j, ++pair_idx) {

            const uint32_t second_path_idx = plan->marginal_posteriors.at(j).second;

            double log_likelihood = loglik[pair_idx];
            log_likelihood += plan->path_log_freqs.at(first_path_idx) + plan->path_log_freqs.at(second_path_idx) + log(numPermutations(vector<uint32_t>({first_path_idx, second_path_idx})));

            if (log_likelihood - max_log_likelihood < plan->min_log_likelihood_diff) { continue; }

            log_likelihoods.emplace_back(log_likelihood);
            max_log_likelihood = max(max_log_likelihood, log_likelihood);
            path_cluster_estimates->path_group_sets.emplace_back(vector<uint32_t>({first_path_idx, second_path_idx}));
        }
    }

    double sum_log_posterior = numeric_limits<double>::lowest();

    for (size_t i = 0; i < log_likelihoods.size(); ++i) {

        if (log_likelihoods.at(i) - max_log_likelihood < plan->min_log_likelihood_diff) { log_likelihoods.at(i) = numeric_limits<double>::lowest(); }
        sum_log_posterior = add_log(sum_log_posterior, log_likelihoods.at(i));
    }

    for (auto & ll: log_likelihoods) { path_cluster_estimates->posteriors.emplace_back(exp(ll - sum_log_posterior)); }

    plan->stage = 2;
    return false;
}

void PathEstimator::finishBoundedPathGroupPosteriors(GroupPosteriorPlan * plan, PathClusterEstimates * path_cluster_estimates, const size_t n_candidates, const uint32_t * cand_pair, const double * cand_loglik) const {

    assert(plan->bounded && plan->stage == 1);

    const uint32_t G = plan->marginal_posteriors.size();

    vector<double> log_likelihoods;
    double max_log_likelihood = numeric_limits<double>::lowest();

    // pair index -> (i, j): row i of the enumeration starts at i * G - i (i - 1) / 2; the candidates come ascending
    uint32_t i = 0;
    uint64_t row_start = 0;

    for (size_t k = 0; k < n_candidates; ++k) {

        const uint64_t pair_idx = cand_pair[k];
        while (i + 1 < G && row_start + (G - i) <= pair_idx) {
            row_start += G - i;
            ++i;
        }

        const uint32_t j = i + static_cast<uint32_t>(pair_idx - row_start);
        assert(j < G);

        const uint32_t first_path_idx = plan->marginal_posteriors.at(i).second;
        const uint32_t second_path_idx = plan->marginal_posteriors.at(j).second;

        double log_likelihood = cand_loglik[k];
        log_likelihood += plan->path_log_freqs.at(first_path_idx) + plan->path_log_freqs.at(second_path_idx) + log(numPermutations(vector<uint32_t>({first_path_idx, second_path_idx})));

        if (log_likelihood - max_log_likelihood < plan->min_log_likelihood_diff) { continue; }

        log_likelihoods.emplace_back(log_likelihood);
        max_log_likelihood = max(max_log_likelihood, log_likelihood);
        path_cluster_estimates->path_group_sets.emplace_back(vector<uint32_t>({first_path_idx, second_path_idx}));
    }

    double sum_log_posterior = numeric_limits<double>::lowest();

    for (size_t k = 0; k < log_likelihoods.size(); ++k) {

        if (log_likelihoods.at(k) - max_log_likelihood < plan->min_log_likelihood_diff) { log_likelihoods.at(k) = numeric_limits<double>::lowest(); }
        sum_log_posterior = add_log(sum_log_posterior, log_likelihoods.at(k));
    }

    for (auto & ll: log_likelihoods) { path_cluster_estimates->posteriors.emplace_back(exp(ll - sum_log_posterior)); }

    plan->stage = 2;
}

void PathEstimator::calculatePathGroupPosteriorsFull(PathClusterEstimates * path_cluster_estimates, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size) const {

    assert(!noise.empty());

    GroupPosteriorPlan plan;
    planPathGroupPosteriors(&plan, path_cluster_estimates, num_paths, path_counts, group_size, false, 1);

    auto loglik = deviceLogLikelihoods(threadContext(), matrix, noise, counts, num_paths, plan.set_size, plan.flat_sets);
    advancePathGroupPosteriors(&plan, path_cluster_estimates, loglik.data());
}

void PathEstimator::calculatePathGroupPosteriorsBounded(PathClusterEstimates * path_cluster_estimates, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size, const double min_rel_likelihood) const {

    assert(group_size == 2);

    // the matrix goes up once; marginals, then the pair walk on the device (the same two rounds the batched drivers run)
    vector<GroupPosteriorPlan> plans(1);
    planPathGroupPosteriors(&plans[0], path_cluster_estimates, num_paths, path_counts, group_size, true, min_rel_likelihood);

    auto * ctx = threadContext();
    const ResidentCluster resident(ctx, matrix, noise, counts, num_paths);

    runPosteriorRounds(ctx, resident.handle, {0, num_paths}, &plans, {path_cluster_estimates});
}

void PathEstimator::estimatePathGroupPosteriorsGibbs(PathClusterEstimates * path_cluster_estimates, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size, mt19937 * mt_rng) const {

    assert(group_size > 0);
    assert(group_size <= 8);

    auto path_log_freqs = calcPathLogFrequences(path_counts);

    path_cluster_estimates->resetEstimates(0, 0);

    uniform_int_distribution<uint32_t> init_path_sampler(0, num_paths - 1);

    // Neither map's iteration order is observable (one is a cache, the other a first-seen index).
    rpvg_b200_host::hash_map<vector<uint32_t>, discrete_distribution<uint32_t> > group_path_sampler_cache;
    rpvg_b200_host::hash_map<vector<uint32_t>, uint32_t> path_group_sets_indices;

    vector<uint32_t> path_group_sample_counts;

    const uint32_t num_gibbs_chains = 10 + round(0.01 * group_size * num_paths);
    const uint32_t num_burn_its = 50 + round(0.025 * group_size * num_paths);
    const uint32_t num_gibbs_its = 100 + round(0.05 * group_size * num_paths);

    auto * ctx = threadContext();
    const ResidentCluster resident(ctx, matrix, noise, counts, num_paths);   // the matrix goes up once for the whole chain

    // Conditional of one slot given the others: log-likelihood of {others} + {k} for every path k. The device adds
    // noise + col / group_size over the set in the order given: the other slots first, the candidate last, like the
    // reference does.
    auto conditional = [&](const vector<uint32_t> & others) {

        vector<uint32_t> flat_sets;
        flat_sets.reserve(static_cast<size_t>(num_paths) * group_size);

        for (uint32_t k = 0; k < num_paths; ++k) {
            flat_sets.insert(flat_sets.end(), others.begin(), others.end());
            flat_sets.push_back(k);
        }

        auto group_probs = resident.logLikelihoods(group_size, flat_sets);

        double sum_log_group_probs = numeric_limits<double>::lowest();

        for (uint32_t k = 0; k < num_paths; ++k) {
            group_probs.at(k) += path_log_freqs.at(k);
            sum_log_group_probs = add_log(sum_log_group_probs, group_probs.at(k));
        }

        for (auto & prob: group_probs) { prob = exp(prob - sum_log_group_probs); }

        return discrete_distribution<uint32_t>(group_probs.begin(), group_probs.end());
    };

    vector<uint32_t> cur_sampled_group_paths;
    cur_sampled_group_paths.reserve(group_size);

    for (uint32_t c = 0; c < num_gibbs_chains; ++c) {

        cur_sampled_group_paths.clear();
        for (uint32_t i = 0; i < group_size; ++i) { cur_sampled_group_paths.emplace_back(init_path_sampler(*mt_rng)); }

        for (uint32_t i = 0; i < num_burn_its + num_gibbs_its; ++i) {

            for (uint32_t j = 0; j < group_size; ++j) {

                vector<uint32_t> key = cur_sampled_group_paths;
                key.at(j) = num_paths;
                sort(key.begin(), key.end());

                auto cache_it = group_path_sampler_cache.emplace(key, discrete_distribution<uint32_t>());

                if (cache_it.second) {

                    vector<uint32_t> others;
                    for (uint32_t k = 0; k < group_size; ++k) {
                        if (j != k) { others.emplace_back(cur_sampled_group_paths.at(k)); }
                    }
                    cache_it.first->second = conditional(others);
                }

                cur_sampled_group_paths.at(j) = cache_it.first->second(*mt_rng);
            }

            if (i >= num_burn_its) {

                vector<uint32_t> sorted_paths = cur_sampled_group_paths;
                sort(sorted_paths.begin(), sorted_paths.end());

                auto index_it = path_group_sets_indices.emplace(sorted_paths, path_cluster_estimates->path_group_sets.size());

                if (index_it.second) {
                    path_cluster_estimates->path_group_sets.emplace_back(sorted_paths);
                    path_group_sample_counts.emplace_back(1);
                } else {
                    path_group_sample_counts.at(index_it.first->second)++;
                }
            }
        }
    }

    for (auto & sample_count: path_group_sample_counts) {
        path_cluster_estimates->posteriors.emplace_back(sample_count / static_cast<double>(num_gibbs_chains * num_gibbs_its));
    }
}

// ---- transcripts -----------------------------------------------------------------------------------------------------

PathAbundanceEstimator::PathAbundanceEstimator(const uint32_t max_em_its_in, const double max_rel_em_conv_in, const uint32_t num_gibbs_samples_in, const uint32_t gibbs_thin_its_in, const double prob_precision) : PathEstimator(prob_precision), max_em_its(max_em_its_in), max_rel_em_conv(max_rel_em_conv_in), num_gibbs_samples(num_gibbs_samples_in), gibbs_thin_its(gibbs_thin_its_in) {}

void PathAbundanceEstimator::estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) {

    path_cluster_estimates->resetEstimates(path_cluster_estimates->paths.size(), 1);

    if (cluster_probs.empty()) { return; }

    const FlatCluster flat(cluster_probs, path_cluster_estimates->paths.size());
    rpvg_b200_clusters clusters = flat.view();

    rpvg_b200_em_params params;
    rpvg_b200_em_params_default(&params);
    params.max_em_its = max_em_its;
    params.max_rel_em_conv = max_rel_em_conv;

    auto * ctx = threadContext();

    // Gibbs samples (rpvg:src/path_abundance_estimator.cpp:34-43): the device sampler is counter-based; its key is drawn
    // from the caller's generator, which keeps the cluster's result a function of (rng_seed + cluster rank)
    const uint32_t num_paths = path_cluster_estimates->paths.size();
    uint64_t gibbs_seed = 0;
    if (num_gibbs_samples > 0) { gibbs_seed = (static_cast<uint64_t>((*mt_rng)()) << 32) | (*mt_rng)(); }

    CountSamples samples;
    rpvg_b200_gibbs gibbs = {&num_gibbs_samples, gibbs_thin_its, &gibbs_seed, nullptr, nullptr};

    if (num_gibbs_samples > 0) {

        samples.path_ids.resize(num_paths);
        iota(samples.path_ids.begin(), samples.path_ids.end(), 0);
        samples.noise_samples.resize(num_gibbs_samples);
        samples.abundance_samples.resize(static_cast<size_t>(num_gibbs_samples) * num_paths);

        gibbs.abundance_samples = samples.abundance_samples.data();
        gibbs.noise_samples = samples.noise_samples.data();
    }

    if (rpvg_b200_em_abundances_gibbs(ctx, &clusters, &params, num_gibbs_samples > 0 ? &gibbs : nullptr, path_cluster_estimates->abundances.data(),
                                      &path_cluster_estimates->noise_count, &path_cluster_estimates->total_count, nullptr) != RPVG_B200_OK) {
        fatal(ctx, "rpvg_b200_em_abundances");
    }

    if (num_gibbs_samples > 0) { path_cluster_estimates->gibbs_read_count_samples.emplace_back(std::move(samples)); }
}

MinimumPathAbundanceEstimator::MinimumPathAbundanceEstimator(const uint32_t max_em_its, const double max_rel_em_conv, const uint32_t num_gibbs_samples, const uint32_t gibbs_thin_its, const double prob_precision) : PathAbundanceEstimator(max_em_its, max_rel_em_conv, num_gibbs_samples, gibbs_thin_its, prob_precision) {}

void MinimumPathAbundanceEstimator::estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) {

    const uint32_t num_paths = path_cluster_estimates->paths.size();
    path_cluster_estimates->resetEstimates(num_paths, 1);

    if (cluster_probs.empty()) { return; }

    const FlatCluster flat(cluster_probs, num_paths);
    rpvg_b200_clusters clusters = flat.view();

    auto * ctx = threadContext();

    // rpvg:src/path_abundance_estimator.cpp:229-252
    vector<uint8_t> in_cover(num_paths, 0);
    if (rpvg_b200_path_cover(ctx, &clusters, in_cover.data()) != RPVG_B200_OK) { fatal(ctx, "rpvg_b200_path_cover"); }

    vector<uint32_t> min_path_cover;
    for (uint32_t j = 0; j < num_paths; ++j) {
        if (in_cover[j]) { min_path_cover.push_back(j); }
    }

    if (min_path_cover.empty()) { return; }

    // :254-269 partial matrix on the cover, normalise, collapse, EM
    const uint32_t subset_cluster = 0;
    const uint64_t subset_offset[2] = {0, min_path_cover.size()};
    rpvg_b200_subsets subsets = {1, &subset_cluster, subset_offset, min_path_cover.data()};

    rpvg_b200_em_params params;
    rpvg_b200_em_params_default(&params);
    params.max_em_its = max_em_its;
    params.max_rel_em_conv = max_rel_em_conv;

    vector<double> cover_abundances(min_path_cover.size(), 0);
    double noise_count = 0, total_count = 0;

    // :271-282 Gibbs samples on the cover's collapsed matrix
    uint64_t gibbs_seed = 0;
    if (num_gibbs_samples > 0) { gibbs_seed = (static_cast<uint64_t>((*mt_rng)()) << 32) | (*mt_rng)(); }

    CountSamples samples;
    rpvg_b200_gibbs gibbs = {&num_gibbs_samples, gibbs_thin_its, &gibbs_seed, nullptr, nullptr};

    if (num_gibbs_samples > 0) {

        samples.path_ids = min_path_cover;
        samples.noise_samples.resize(num_gibbs_samples);
        samples.abundance_samples.resize(static_cast<size_t>(num_gibbs_samples) * min_path_cover.size());

        gibbs.abundance_samples = samples.abundance_samples.data();
        gibbs.noise_samples = samples.noise_samples.data();
    }

    if (rpvg_b200_subset_em_gibbs(ctx, &clusters, &subsets, &params, prob_precision, num_gibbs_samples > 0 ? &gibbs : nullptr, cover_abundances.data(), &noise_count,
                                  &total_count, nullptr, nullptr) != RPVG_B200_OK) {
        fatal(ctx, "rpvg_b200_subset_em");
    }

    if (num_gibbs_samples > 0) { path_cluster_estimates->gibbs_read_count_samples.emplace_back(std::move(samples)); }

    // :286-292
    for (size_t i = 0; i < min_path_cover.size(); ++i) { path_cluster_estimates->abundances.at(min_path_cover.at(i)) += cover_abundances.at(i); }

    path_cluster_estimates->noise_count = noise_count;
    path_cluster_estimates->total_count = total_count;
}

// rpvg:src/path_abundance_estimator.cpp:297-340, the greedy picks on the device (rpvg_b200_weighted_path_cover)
vector<uint32_t> MinimumPathAbundanceEstimator::weightedMinimumPathCover(const uint32_t rows, const uint32_t n_paths, const uint8_t * read_path_cover, const double * read_counts, const double * path_weights) const {

    vector<uint64_t> row_ptr = {0};
    vector<uint32_t> col;

    for (uint32_t i = 0; i < rows; ++i) {
        for (uint32_t j = 0; j < n_paths; ++j) {
            if (read_path_cover[static_cast<size_t>(j) * rows + i]) { col.push_back(j); }
        }
        row_ptr.push_back(col.size());
    }

    auto * ctx = threadContext();
    vector<uint8_t> in_cover(n_paths, 0);

    if (rpvg_b200_weighted_path_cover(ctx, rows, n_paths, row_ptr.data(), col.data(), read_counts, path_weights, in_cover.data()) != RPVG_B200_OK) {
        fatal(ctx, "rpvg_b200_weighted_path_cover");
    }

    vector<uint32_t> min_path_cover;
    for (uint32_t j = 0; j < n_paths; ++j) {
        if (in_cover[j]) { min_path_cover.push_back(j); }
    }

    return min_path_cover;
}

#ifdef RPVG_B200_WITH_RPVG_HEADERS
vector<uint32_t> MinimumPathAbundanceEstimator::weightedMinimumPathCover(const Utils::ColMatrixXb & read_path_cover, const Utils::RowVectorXd & read_counts, const Utils::RowVectorXd & path_weights) const {

    assert(read_path_cover.rows() == read_counts.cols());
    assert(read_path_cover.cols() == path_weights.cols());

    vector<uint8_t> cover(static_cast<size_t>(read_path_cover.rows()) * read_path_cover.cols());
    for (Eigen::Index j = 0; j < read_path_cover.cols(); ++j) {
        for (Eigen::Index i = 0; i < read_path_cover.rows(); ++i) { cover[static_cast<size_t>(j) * read_path_cover.rows() + i] = read_path_cover(i, j) ? 1 : 0; }
    }

    return weightedMinimumPathCover(read_path_cover.rows(), read_path_cover.cols(), cover.data(), read_counts.data(), path_weights.data());
}
#endif

// ---- haplotypes ------------------------------------------------------------------------------------------------------

static const double min_rel_likelihood = 1e-8;   // rpvg:src/path_posterior_estimator.cpp:5

PathPosteriorEstimator::PathPosteriorEstimator(const double prob_precision) : PathEstimator(prob_precision) {}

void PathPosteriorEstimator::estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) {

    path_cluster_estimates->resetEstimates(path_cluster_estimates->paths.size(), 1);

    if (cluster_probs.empty()) { return; }

    vector<double> matrix, noise, counts;
    denseMatrix(&matrix, &noise, &counts, cluster_probs, path_cluster_estimates->paths.size());

    vector<uint32_t> path_counts;
    for (auto & path: path_cluster_estimates->paths) { path_counts.emplace_back(path.source_count); }

    calculatePathGroupPosteriorsFull(path_cluster_estimates, matrix, noise, counts, path_cluster_estimates->paths.size(), path_counts, 1);
}

PathGroupPosteriorEstimator::PathGroupPosteriorEstimator(const uint32_t group_size_in, const bool use_group_post_gibbs_in, const double prob_precision) : PathPosteriorEstimator(prob_precision), group_size(group_size_in), use_group_post_gibbs(use_group_post_gibbs_in) {}

void PathGroupPosteriorEstimator::estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) {

    path_cluster_estimates->resetEstimates(0, 0);

    if (cluster_probs.empty()) { return; }

    const uint32_t num_paths = path_cluster_estimates->paths.size();

    vector<double> matrix, noise, counts;
    denseMatrix(&matrix, &noise, &counts, cluster_probs, num_paths);

    vector<uint32_t> path_counts;
    for (auto & path: path_cluster_estimates->paths) { path_counts.emplace_back(path.source_count); }

    if (use_group_post_gibbs) {

        estimatePathGroupPosteriorsGibbs(path_cluster_estimates, matrix, noise, counts, num_paths, path_counts, group_size, mt_rng);

    } else if (group_size == 2) {

        calculatePathGroupPosteriorsBounded(path_cluster_estimates, matrix, noise, counts, num_paths, path_counts, group_size, min_rel_likelihood);

    } else {

        calculatePathGroupPosteriorsFull(path_cluster_estimates, matrix, noise, counts, num_paths, path_counts, group_size);
    }
}
