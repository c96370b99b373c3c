// Batched drivers: the cluster loop of rpvg:src/main.cpp:829-977 re-stated as "flatten all clusters, a few device calls
// for the whole batch, scatter the results" (SURVEY.md §8(f) rank 1). Per cluster the device calls of estimate() cost a
// few hundred microseconds of launches, copies and synchronisation whatever the cluster's size; clusters of a real run
// are mostly small, so the per-cluster drop-in is launch-bound (tools/nested_bench.py). The batch entry points give the
// same results cluster by cluster: they run the same staged host code around the same C ABI calls, which take many
// clusters as easily as one.

#include "estimators.hpp"

#include <algorithm>
#include <cassert>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <numeric>
#include <thread>

#include <omp.h>

#include "rpvg_b200.h"

namespace {

[[noreturn]] void fatalBatch(rpvg_b200_ctx * ctx, const char * what) {

    cerr << "ERROR: rpvg_b200 " << what << ": " << rpvg_b200_last_error(ctx) << endl;
    abort();
}

// Many clusters as one arena (include/rpvg_b200.h): every pair<double, vector<uint32_t>> expanded, entries of a read
// class sorted by path — the flattening estimate() does for one cluster.
struct BatchArena {

    vector<uint64_t> read_offset = {0}, path_offset = {0};

    // filled completely by build(), in page-locked memory of the thread's context (rpvg_b200_pinned_scratch: the
    // arena goes to the device right away, twice for the nested model, and pageable memory would halve that rate)
    uint64_t * row_ptr = nullptr;
    uint32_t * col = nullptr, * count = nullptr;
    double * prob = nullptr, * noise = nullptr;

    // all clusters at once, filled in parallel over stretches of read classes (a cluster may hold most of the batch: the
    // unit of work is 256 read classes, not a cluster)
    // (sorted = false: the entries of a read class stay in the order of its groups — enough for the dense matrices of
    // the haplotypes model, which are filled by a scatter)
    void build(rpvg_b200_ctx * ctx, const vector<const vector<ReadPathProbabilities> *> & members, const vector<uint32_t> & num_paths, const bool sorted = true) {

        const size_t n = members.size();

        read_offset.assign(1, 0);
        path_offset.assign(1, 0);

        for (size_t k = 0; k < n; ++k) {
            read_offset.push_back(read_offset.back() + members[k]->size());
            path_offset.push_back(path_offset.back() + num_paths[k]);
        }

        const uint64_t R = read_offset.back();

        struct Stretch { uint32_t cluster; uint32_t begin, end; };
        constexpr uint32_t kStretchRows = 256;

        vector<Stretch> stretches;
        for (size_t k = 0; k < n; ++k) {
            const uint32_t m = members[k]->size();
            for (uint32_t b = 0; b < m; b += kStretchRows) { stretches.push_back({static_cast<uint32_t>(k), b, min(m, b + kStretchRows)}); }
        }

        row_ptr = static_cast<uint64_t *>(rpvg_b200_pinned_scratch(ctx, 0, sizeof(uint64_t) * (R + 1)));
        noise = static_cast<double *>(rpvg_b200_pinned_scratch(ctx, 3, sizeof(double) * (R + 1)));
        count = static_cast<uint32_t *>(rpvg_b200_pinned_scratch(ctx, 4, sizeof(uint32_t) * (R + 1)));
        if (!row_ptr || !noise || !count) { fatalBatch(ctx, "rpvg_b200_pinned_scratch"); }

        // pass 1: entries of every read class (into row_ptr[r + 1]), its noise and count
        #pragma omp parallel for schedule(dynamic, 4)
        for (size_t q = 0; q < stretches.size(); ++q) {

            const auto & st = stretches[q];
            const auto & probs = *members[st.cluster];
            uint64_t r = read_offset[st.cluster] + st.begin;

            for (uint32_t i = st.begin; i < st.end; ++i, ++r) {
                uint64_t e = 0;
                for (auto & path_probs: probs[i].pathProbs()) { e += path_probs.second.size(); }
                row_ptr[r + 1] = e;
                noise[r] = probs[i].noiseProb();
                count[r] = probs[i].readCount();
            }
        }

        row_ptr[0] = 0;
        for (uint64_t r = 0; r < R; ++r) { row_ptr[r + 1] += row_ptr[r]; }

        const uint64_t Z = row_ptr[R];

        col = static_cast<uint32_t *>(rpvg_b200_pinned_scratch(ctx, 1, sizeof(uint32_t) * (Z + 1)));
        prob = static_cast<double *>(rpvg_b200_pinned_scratch(ctx, 2, sizeof(double) * (Z + 1)));
        if (!col || !prob) { fatalBatch(ctx, "rpvg_b200_pinned_scratch"); }

        // pass 2: the entries, ascending by path within a read class. A read class comes as a few groups of ascending ids:
        // one group is copied as it is, several are merged by a sort
        #pragma omp parallel for schedule(dynamic, 4)
        for (size_t q = 0; q < stretches.size(); ++q) {

            const auto & st = stretches[q];
            const auto & probs = *members[st.cluster];
            uint64_t r = read_offset[st.cluster] + st.begin;

            vector<pair<uint32_t, double> > row;

            for (uint32_t i = st.begin; i < st.end; ++i, ++r) {

                uint64_t e = row_ptr[r];
                const auto & groups = probs[i].pathProbs();

                if (!sorted || (groups.size() == 1 && is_sorted(groups.front().second.begin(), groups.front().second.end()))) {

                    for (auto & path_probs: groups) {
                        const double p = path_probs.first;
                        for (auto & path: path_probs.second) {
                            col[e] = path;
                            prob[e] = p;
                            ++e;
                        }
                    }
                    continue;
                }

                // several groups, each ascending (the usual case): merged by always taking the smallest head
                if (groups.size() <= 8) {

                    const uint32_t * head[8];
                    const uint32_t * tail[8];
                    double value[8];
                    size_t k = 0;
                    bool ascending = true;

                    for (auto & path_probs: groups) {
                        ascending = ascending && is_sorted(path_probs.second.begin(), path_probs.second.end());
                        if (path_probs.second.empty()) { continue; }
                        head[k] = path_probs.second.data();
                        tail[k] = head[k] + path_probs.second.size();
                        value[k] = path_probs.first;
                        ++k;
                    }

                    if (ascending) {

                        while (k > 0) {

                            size_t best = 0;
                            for (size_t g = 1; g < k; ++g) {
                                if (*head[g] < *head[best]) { best = g; }
                            }

                            col[e] = *head[best];
                            prob[e] = value[best];
                            ++e;

                            if (++head[best] == tail[best]) {
                                --k;
                                head[best] = head[k];
                                tail[best] = tail[k];
                                value[best] = value[k];
                            }
                        }
                        continue;
                    }
                }

                row.clear();
                for (auto & path_probs: groups) {
                    for (auto & path: path_probs.second) { row.emplace_back(path, path_probs.first); }
                }
                sort(row.begin(), row.end());

                for (auto & entry: row) {
                    col[e] = entry.first;
                    prob[e] = entry.second;
                    ++e;
                }
            }
        }
    }

    uint32_t size() const { return static_cast<uint32_t>(read_offset.size() - 1); }

    rpvg_b200_clusters view() const {
        return rpvg_b200_clusters{size(), read_offset.data(), path_offset.data(), row_ptr, col, prob, noise, count};
    }
};

}

namespace {

// Many clusters as read records (rpvg_b200_read_records): every pair<double, vector<uint32_t>> stays a probability and its
// list of paths — what the dense matrices of the haplotypes model are filled from. A third of the arena's bytes.
struct BatchRecords {

    vector<uint64_t> read_offset = {0}, path_offset = {0};

    // in page-locked memory of the thread's context (rpvg_b200_pinned_scratch)
    uint64_t * group_ptr = nullptr, * id_ptr = nullptr;
    uint32_t * ids = nullptr, * count = nullptr;
    double * group_prob = nullptr, * noise = nullptr;

    void build(rpvg_b200_ctx * ctx, const vector<const vector<ReadPathProbabilities> *> & members, const vector<uint32_t> & num_paths) {

        const size_t n = members.size();

        read_offset.assign(1, 0);
        path_offset.assign(1, 0);

        for (size_t k = 0; k < n; ++k) {
            read_offset.push_back(read_offset.back() + members[k]->size());
            path_offset.push_back(path_offset.back() + num_paths[k]);
        }

        const uint64_t R = read_offset.back();

        struct Stretch { uint32_t cluster; uint32_t begin, end; };
        constexpr uint32_t kStretchRows = 256;

        vector<Stretch> stretches;
        for (size_t k = 0; k < n; ++k) {
            const uint32_t m = members[k]->size();
            for (uint32_t b = 0; b < m; b += kStretchRows) { stretches.push_back({static_cast<uint32_t>(k), b, min(m, b + kStretchRows)}); }
        }

        group_ptr = static_cast<uint64_t *>(rpvg_b200_pinned_scratch(ctx, 0, sizeof(uint64_t) * (R + 1)));
        noise = static_cast<double *>(rpvg_b200_pinned_scratch(ctx, 3, sizeof(double) * (R + 1)));
        count = static_cast<uint32_t *>(rpvg_b200_pinned_scratch(ctx, 4, sizeof(uint32_t) * (R + 1)));
        if (!group_ptr || !noise || !count) { fatalBatch(ctx, "rpvg_b200_pinned_scratch"); }

        vector<uint64_t> first_id(R + 1, 0);

        // pass 1: groups and ids of every record
        #pragma omp parallel for schedule(dynamic, 4)
        for (size_t q = 0; q < stretches.size(); ++q) {

            const auto & st = stretches[q];
            const auto & probs = *members[st.cluster];
            uint64_t r = read_offset[st.cluster] + st.begin;

            for (uint32_t i = st.begin; i < st.end; ++i, ++r) {
                uint64_t e = 0;
                for (auto & path_probs: probs[i].pathProbs()) { e += path_probs.second.size(); }
                group_ptr[r + 1] = probs[i].pathProbs().size();
                first_id[r + 1] = e;
                noise[r] = probs[i].noiseProb();
                count[r] = probs[i].readCount();
            }
        }

        group_ptr[0] = 0;
        for (uint64_t r = 0; r < R; ++r) {
            group_ptr[r + 1] += group_ptr[r];
            first_id[r + 1] += first_id[r];
        }

        const uint64_t NG = group_ptr[R], Z = first_id[R];

        group_prob = static_cast<double *>(rpvg_b200_pinned_scratch(ctx, 2, sizeof(double) * (NG + 1)));
        id_ptr = static_cast<uint64_t *>(rpvg_b200_pinned_scratch(ctx, 5, sizeof(uint64_t) * (NG + 1)));
        ids = static_cast<uint32_t *>(rpvg_b200_pinned_scratch(ctx, 1, sizeof(uint32_t) * (Z + 1)));
        if (!group_prob || !id_ptr || !ids) { fatalBatch(ctx, "rpvg_b200_pinned_scratch"); }

        // pass 2: the groups as they are
        #pragma omp parallel for schedule(dynamic, 4)
        for (size_t q = 0; q < stretches.size(); ++q) {

            const auto & st = stretches[q];
            const auto & probs = *members[st.cluster];
            uint64_t r = read_offset[st.cluster] + st.begin;

            for (uint32_t i = st.begin; i < st.end; ++i, ++r) {

                uint64_t g = group_ptr[r];
                uint64_t e = first_id[r];

                for (auto & path_probs: probs[i].pathProbs()) {
                    group_prob[g] = path_probs.first;
                    id_ptr[g] = e;
                    if (!path_probs.second.empty()) { memcpy(ids + e, path_probs.second.data(), sizeof(uint32_t) * path_probs.second.size()); }
                    e += path_probs.second.size();
                    ++g;
                }
            }
        }

        id_ptr[NG] = Z;
    }

    rpvg_b200_read_records view() const {
        return rpvg_b200_read_records{static_cast<uint32_t>(read_offset.size() - 1), read_offset.data(), group_ptr, group_prob, id_ptr, ids, noise, count};
    }
};

// stage times on stderr when RPVG_B200_BATCH_TRACE is set
struct BatchTrace {

    const bool on = getenv("RPVG_B200_BATCH_TRACE") != nullptr;
    chrono::steady_clock::time_point last = chrono::steady_clock::now();

    void lap(const char * what) {
        if (!on) { return; }
        const auto now = chrono::steady_clock::now();
        cerr << "[rpvg_b200 batch] " << what << ": " << chrono::duration<double, milli>(now - last).count() << " ms" << endl;
        last = now;
    }
};

}

void PathEstimator::runPosteriorRounds(rpvg_b200_ctx * ctx, rpvg_b200_grouped * resident, const vector<uint64_t> & group_offset, vector<GroupPosteriorPlan> * plans, const vector<PathClusterEstimates *> & outs) const {

    const size_t n = plans->size();
    if (n == 0) { return; }

    const bool bounded = plans->front().bounded;
    for (auto & plan: *plans) { plan.device_pairs = bounded; }

    BatchTrace trace;

    // round 1: the first request of every plan (Full: all multisets; Bounded: the single groups)
    {
        const uint32_t set_size = plans->front().set_size;

        vector<uint64_t> set_offset = {0};
        for (auto & plan: *plans) {
            assert(plan.set_size == set_size && plan.bounded == bounded);
            set_offset.push_back(set_offset.back() + plan.flat_sets.size() / set_size);
        }

        vector<uint32_t> sets(set_offset.back() * set_size);

        #pragma omp parallel for schedule(dynamic, 8)
        for (size_t k = 0; k < n; ++k) { copy((*plans)[k].flat_sets.begin(), (*plans)[k].flat_sets.end(), sets.begin() + set_offset[k] * set_size); }

        vector<double> loglik(set_offset.back(), 0);
        if (rpvg_b200_grouped_loglik(ctx, resident, set_size, set_offset.data(), sets.data(), loglik.data()) != RPVG_B200_OK) { fatalBatch(ctx, "rpvg_b200_grouped_loglik"); }

        #pragma omp parallel for schedule(dynamic, 8)
        for (size_t k = 0; k < n; ++k) {
            const bool more = advancePathGroupPosteriors(&(*plans)[k], outs[k], loglik.data() + set_offset[k]);
            assert(more == bounded);
            (void) more;
        }
    }

    trace.lap("posteriors: round 1 (marginals / all sets)");

    if (!bounded) { return; }

    // round 2 (Bounded): the pairs are enumerated, evaluated and pre-filtered on the device from each cluster's order of
    // groups; the reference's walk is replayed over the candidates (a margin of 1e-6 in log space keeps every pair the
    // exact test could keep)
    vector<uint32_t> order(group_offset.back());
    vector<double> log_freq(group_offset.back());

    #pragma omp parallel for schedule(dynamic, 8)
    for (size_t k = 0; k < n; ++k) {
        const auto & plan = (*plans)[k];
        assert(plan.marginal_posteriors.size() == group_offset[k + 1] - group_offset[k]);
        for (size_t i = 0; i < plan.marginal_posteriors.size(); ++i) {
            order[group_offset[k] + i] = plan.marginal_posteriors[i].second;
            log_freq[group_offset[k] + i] = plan.path_log_freqs[i];
        }
    }

    const uint64_t * cand_offset = nullptr;
    const uint32_t * cand_pair = nullptr;
    const double * cand_loglik = nullptr;

    if (rpvg_b200_grouped_bounded_pairs(ctx, resident, group_offset.data(), order.data(), log_freq.data(), plans->front().min_log_likelihood_diff - 1e-6, &cand_offset, &cand_pair,
                                        &cand_loglik) != RPVG_B200_OK) {
        fatalBatch(ctx, "rpvg_b200_grouped_bounded_pairs");
    }

    trace.lap("posteriors: pairs on the device");

    #pragma omp parallel for schedule(dynamic, 8)
    for (size_t k = 0; k < n; ++k) {
        const uint64_t c0 = cand_offset[k];
        finishBoundedPathGroupPosteriors(&(*plans)[k], outs[k], cand_offset[k + 1] - c0, cand_pair ? cand_pair + c0 : nullptr, cand_loglik ? cand_loglik + c0 : nullptr);
    }

    trace.lap("posteriors: walk over the candidates");
}

void PathEstimator::estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs) {

    assert(estimates.size() == cluster_probs.size());
    assert(estimates.size() == rngs.size());

    for (size_t i = 0; i < estimates.size(); ++i) { estimateOne(estimates[i], *cluster_probs[i], rngs[i]); }
}

// One batch over the GPUs of the box: see estimators.hpp. The split mirrors the reference's own schedule — largest first,
// each to the least loaded worker (rpvg:src/main.cpp:827-830) — with devices in place of threads.
void PathEstimator::estimateBatchOnDevices(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs, int num_devices) {

    assert(estimates.size() == cluster_probs.size());
    assert(estimates.size() == rngs.size());

    if (num_devices <= 0) {
        const char * env = getenv("RPVG_B200_DEVICES");
        num_devices = 1;
        if (env && string(env) == "all") { num_devices = rpvg_b200_device_count(); } else if (env) { num_devices = atoi(env); }
    }

    num_devices = max(1, min(num_devices, rpvg_b200_device_count()));
    num_devices = static_cast<int>(min<size_t>(num_devices, max<size_t>(1, estimates.size())));

    if (num_devices == 1) {
        estimateBatch(estimates, cluster_probs, rngs);
        return;
    }

    const size_t n = estimates.size();
    vector<double> cost(n, 0);

    #pragma omp parallel for schedule(dynamic, 16)
    for (size_t i = 0; i < n; ++i) { cost[i] = clusterCost(*estimates[i], *cluster_probs[i]); }

    // longest processing time first; ties by index, so the split is deterministic
    vector<size_t> order(n);
    iota(order.begin(), order.end(), 0);
    stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return cost[a] > cost[b]; });

    vector<vector<size_t> > share(num_devices);
    vector<double> load(num_devices, 0);

    for (auto i: order) {
        const size_t d = min_element(load.begin(), load.end()) - load.begin();
        share[d].push_back(i);
        load[d] += cost[i];
    }

    const int first_device = threadDevice();
    const int host_threads = max(1, omp_get_max_threads() / num_devices);

    vector<thread> workers;

    for (int d = 0; d < num_devices; ++d) {

        workers.emplace_back([&, d]() {

            sort(share[d].begin(), share[d].end());   // the caller's order within a share

            vector<PathClusterEstimates *> est;
            vector<const vector<ReadPathProbabilities> *> probs;
            vector<mt19937 *> generators;

            for (auto i: share[d]) {
                est.push_back(estimates[i]);
                probs.push_back(cluster_probs[i]);
                generators.push_back(rngs[i]);
            }

            threadDevice() = (first_device + d) % rpvg_b200_device_count();
            omp_set_num_threads(host_threads);

            if (!est.empty()) { estimateBatch(est, probs, generators); }
        });
    }

    for (auto & worker: workers) { worker.join(); }
}

// transcripts model: EM (and Gibbs read counts) of all clusters in one device batch
void PathAbundanceEstimator::estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs) {

    assert(estimates.size() == cluster_probs.size());
    assert(estimates.size() == rngs.size());

    BatchArena arena;
    vector<size_t> members;   // clusters with read classes, in order

    for (size_t i = 0; i < estimates.size(); ++i) {

        estimates[i]->resetEstimates(estimates[i]->paths.size(), 1);

        if (!cluster_probs[i]->empty()) { members.push_back(i); }
    }

    if (members.empty()) { return; }

    {
        vector<const vector<ReadPathProbabilities> *> member_probs;
        vector<uint32_t> member_paths;
        for (auto i: members) {
            member_probs.push_back(cluster_probs[i]);
            member_paths.push_back(estimates[i]->paths.size());
        }
        arena.build(threadContext(), member_probs, member_paths);
    }

    rpvg_b200_em_params params;
    rpvg_b200_em_params_default(&params);
    params.max_em_its = max_em_its;
    params.max_rel_em_conv = max_rel_em_conv;

    // the sampler's keys come from each cluster's own generator, as in estimate()
    vector<uint32_t> gibbs_counts(members.size(), num_gibbs_samples);
    vector<uint64_t> gibbs_seeds(members.size(), 0);
    size_t total_values = 0;

    if (num_gibbs_samples > 0) {
        for (size_t k = 0; k < members.size(); ++k) {
            mt19937 * mt_rng = rngs[members[k]];
            gibbs_seeds[k] = (static_cast<uint64_t>((*mt_rng)()) << 32) | (*mt_rng)();
            total_values += static_cast<size_t>(num_gibbs_samples) * estimates[members[k]]->paths.size();
        }
    }

    vector<double> abundances(arena.path_offset.back());
    vector<double> noise_counts(members.size()), total_counts(members.size());
    vector<double> gibbs_abundances(total_values), gibbs_noise(static_cast<size_t>(num_gibbs_samples) * members.size());

    rpvg_b200_gibbs gibbs = {gibbs_counts.data(), gibbs_thin_its, gibbs_seeds.data(), gibbs_abundances.data(), gibbs_noise.data()};
    rpvg_b200_clusters clusters = arena.view();

    auto * ctx = threadContext();

    if (rpvg_b200_em_abundances_gibbs(ctx, &clusters, &params, num_gibbs_samples > 0 ? &gibbs : nullptr, abundances.data(), noise_counts.data(), total_counts.data(),
                                      nullptr) != RPVG_B200_OK) {
        fatalBatch(ctx, "rpvg_b200_em_abundances");
    }

    size_t value_pos = 0;

    for (size_t k = 0; k < members.size(); ++k) {

        PathClusterEstimates * est = estimates[members[k]];
        const size_t num_paths = est->paths.size();
        const size_t p0 = arena.path_offset[k];

        copy(abundances.begin() + p0, abundances.begin() + p0 + num_paths, est->abundances.begin());
        est->noise_count = noise_counts[k];
        est->total_count = total_counts[k];

        if (num_gibbs_samples > 0) {

            CountSamples samples;
            samples.path_ids.resize(num_paths);
            iota(samples.path_ids.begin(), samples.path_ids.end(), 0);
            samples.noise_samples.assign(gibbs_noise.begin() + k * num_gibbs_samples, gibbs_noise.begin() + (k + 1) * num_gibbs_samples);
            samples.abundance_samples.assign(gibbs_abundances.begin() + value_pos, gibbs_abundances.begin() + value_pos + static_cast<size_t>(num_gibbs_samples) * num_paths);
            value_pos += static_cast<size_t>(num_gibbs_samples) * num_paths;

            est->gibbs_read_count_samples.emplace_back(std::move(samples));
        }
    }
}

// haplotype-transcripts model, collapsed mode: the stages of inferAbundancesCollapsedGroups (nested.cpp) for all clusters
// at once, each device request of a stage put into one call
void NestedPathAbundanceEstimator::estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs) {

    assert(estimates.size() == cluster_probs.size());
    assert(estimates.size() == rngs.size());

    if (!infer_collapsed || use_group_post_gibbs) {   // host-side sampling between the device calls of a cluster: looped
        PathEstimator::estimateBatch(estimates, cluster_probs, rngs);
        return;
    }

    struct Item {
        size_t index;
        pair<vector<vector<uint32_t> >, vector<uint32_t> > source_groups;
        PathClusterEstimates group_estimates;
        vector<SubsetJob> jobs;
    };

    const bool trace = getenv("RPVG_B200_BATCH_TRACE") != nullptr;
    auto clock_now = chrono::steady_clock::now();
    auto lap = [&](const char * what) {
        if (!trace) { return; }
        const auto now = chrono::steady_clock::now();
        cerr << "[rpvg_b200 batch] " << what << ": " << chrono::duration<double, milli>(now - clock_now).count() << " ms" << endl;
        clock_now = now;
    };

    // ---- stage A: flatten, haplotype groups
    BatchArena arena;
    vector<Item> items;

    for (size_t i = 0; i < estimates.size(); ++i) {

        estimates[i]->resetEstimates(0, 0);
        if (cluster_probs[i]->empty()) { continue; }

        items.emplace_back();
        items.back().index = i;
    }

    if (items.empty()) { return; }

    {
        vector<const vector<ReadPathProbabilities> *> members;
        vector<uint32_t> member_paths;
        for (auto & item: items) {
            members.push_back(cluster_probs[item.index]);
            member_paths.push_back(estimates[item.index]->paths.size());
        }
        arena.build(threadContext(), members, member_paths);
    }

    // (the host stages are independent per cluster: all cores, like the reference's loop; the device calls are made by
    // this thread)
    lap("A flatten");

    // (the cost is the number of haplotype ids a cluster's paths carry: the largest clusters go first, one at a time)
    {
        vector<pair<uint64_t, uint32_t> > by_cost(items.size());
        for (size_t k = 0; k < items.size(); ++k) {
            uint64_t ids = 0;
            for (auto & path: estimates[items[k].index]->paths) { ids += path.source_ids.size(); }
            by_cost[k] = make_pair(ids, static_cast<uint32_t>(k));
        }
        sort(by_cost.begin(), by_cost.end(), [](const pair<uint64_t, uint32_t> & a, const pair<uint64_t, uint32_t> & b) { return a.first > b.first || (a.first == b.first && a.second < b.second); });

        #pragma omp parallel for schedule(dynamic, 1)
        for (size_t q = 0; q < by_cost.size(); ++q) {
            const size_t k = by_cost[q].second;
            items[k].source_groups = findPathSourceGroups(estimates[items[k].index]->paths);
        }
    }

    auto * ctx = threadContext();
    rpvg_b200_clusters clusters = arena.view();

    lap("A flatten + haplotype groups");

    // ---- stage B: grouped + normalised + collapsed matrix of every cluster; they stay on the device for stage C
    rpvg_b200_grouped * grouped = nullptr;
    vector<uint64_t> group_offset = {0};
    vector<GroupPosteriorPlan> plans(items.size());
    vector<PathClusterEstimates *> group_outs(items.size());
    {
        vector<uint64_t> group_path_ptr = {0};
        vector<uint32_t> group_paths;

        for (size_t k = 0; k < items.size(); ++k) {

            for (auto & group: items[k].source_groups.first) {
                group_paths.insert(group_paths.end(), group.begin(), group.end());
                group_path_ptr.push_back(group_paths.size());
            }

            group_offset.push_back(group_path_ptr.size() - 1);
        }

        rpvg_b200_path_groups groups = {group_offset.data(), group_path_ptr.data(), group_paths.data()};
        vector<uint32_t> n_classes(items.size(), 0);

        if (rpvg_b200_grouped_create(ctx, &clusters, &groups, prob_precision, n_classes.data(), &grouped) != RPVG_B200_OK) { fatalBatch(ctx, "rpvg_b200_grouped_create"); }

        #pragma omp parallel for schedule(dynamic, 8)
        for (size_t k = 0; k < items.size(); ++k) {
            assert(n_classes[k] > 0);
            group_outs[k] = &items[k].group_estimates;
            planPathGroupPosteriors(&plans[k], group_outs[k], items[k].source_groups.first.size(), items[k].source_groups.second, group_size, group_size == 2, min_hap_prob);
        }
    }

    lap("B grouped collapse");

    // ---- stage C: posteriors of the haplotype combinations (Full: one log-likelihood call for all clusters; Bounded: the
    // marginals, then the pair walk on the device)
    runPosteriorRounds(ctx, grouped, group_offset, &plans, group_outs);
    rpvg_b200_grouped_free(ctx, grouped);

    lap("C posteriors");

    // ---- stage D: the path subsets of every cluster (host; each cluster draws from its own generator, in the order
    // estimate() would)
    vector<uint32_t> subset_cluster, subset_paths, gibbs_counts;
    vector<uint64_t> subset_offset = {0}, gibbs_seeds;
    vector<size_t> first_subset = {0}, first_path = {0}, first_sample = {0}, first_value = {0};

    #pragma omp parallel for schedule(dynamic, 8)
    for (size_t k = 0; k < items.size(); ++k) {

        auto & item = items[k];

        const SubsetSamples path_subset_samples = collapsedSubsetSamples(item.group_estimates, item.source_groups.first);
        item.jobs = planPathSubsets(estimates[item.index], *cluster_probs[item.index], rngs[item.index], path_subset_samples);
    }

    for (size_t k = 0; k < items.size(); ++k) {

        auto & item = items[k];

        size_t samples = first_sample.back(), values = first_value.back();

        for (auto & job: item.jobs) {
            subset_cluster.push_back(static_cast<uint32_t>(k));
            subset_paths.insert(subset_paths.end(), job.collapsed.begin(), job.collapsed.end());
            subset_offset.push_back(subset_paths.size());
            gibbs_counts.push_back(job.gibbs_samples);
            gibbs_seeds.push_back(job.gibbs_seed);
            samples += job.gibbs_samples;
            values += static_cast<size_t>(job.gibbs_samples) * job.collapsed.size();
        }

        first_subset.push_back(subset_cluster.size());
        first_path.push_back(subset_paths.size());
        first_sample.push_back(samples);
        first_value.push_back(values);
    }

    lap("D subsets");

    // ---- stage E: all sub-EMs (and their Gibbs samples) of all clusters in one call
    vector<double> abundances(subset_paths.size()), noise_counts(subset_cluster.size(), 0), total_counts(subset_cluster.size(), 0);
    vector<double> gibbs_abundances(first_value.back()), gibbs_noise(first_sample.back());

    if (!subset_cluster.empty()) {

        rpvg_b200_subsets subsets = {static_cast<uint32_t>(subset_cluster.size()), subset_cluster.data(), subset_offset.data(), subset_paths.data()};

        rpvg_b200_em_params params;
        rpvg_b200_em_params_default(&params);
        params.max_em_its = max_em_its;
        params.max_rel_em_conv = max_rel_em_conv;

        rpvg_b200_gibbs gibbs = {gibbs_counts.data(), gibbs_thin_its, gibbs_seeds.data(), gibbs_abundances.data(), gibbs_noise.data()};

        if (rpvg_b200_subset_em_gibbs(ctx, &clusters, &subsets, &params, prob_precision, first_sample.back() > 0 ? &gibbs : nullptr, abundances.data(), noise_counts.data(),
                                      total_counts.data(), nullptr, nullptr) != RPVG_B200_OK) {
            fatalBatch(ctx, "rpvg_b200_subset_em");
        }
    }

    lap("E sub-EMs");

    // ---- stage F: weight, file under the transcripts' path tuples (host)
    #pragma omp parallel for schedule(dynamic, 8)
    for (size_t k = 0; k < items.size(); ++k) {
        finishPathSubsets(estimates[items[k].index], items[k].jobs, abundances.data() + first_path[k], noise_counts.data() + first_subset[k],
                          gibbs_abundances.data() + first_value[k], gibbs_noise.data() + first_sample[k]);
    }

    lap("F accumulate");
}

// haplotypes model (posteriors over the raw matrix, rpvg:src/path_posterior_estimator.cpp:9-71): dense matrices of all
// clusters built on all cores and uploaded once, then the rounds of the staged posterior calculation
void PathEstimator::posteriorsBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const uint32_t group_size, const bool bounded, const double min_rel_likelihood) const {

    vector<size_t> members;
    for (size_t i = 0; i < estimates.size(); ++i) {
        if (!cluster_probs[i]->empty()) { members.push_back(i); }
    }

    if (members.empty()) { return; }

    const size_t n = members.size();

    BatchTrace trace;

    BatchRecords arena;
    {
        vector<const vector<ReadPathProbabilities> *> member_probs;
        vector<uint32_t> member_paths;
        for (auto i: members) {
            member_probs.push_back(cluster_probs[i]);
            member_paths.push_back(estimates[i]->paths.size());
        }
        arena.build(threadContext(), member_probs, member_paths);
    }

    trace.lap("haplotypes: flatten");

    const vector<uint64_t> & path_offset = arena.path_offset;
    vector<GroupPosteriorPlan> plans(n);

    #pragma omp parallel for schedule(dynamic, 8)
    for (size_t k = 0; k < n; ++k) {

        PathClusterEstimates * est = estimates[members[k]];

        vector<uint32_t> path_counts;
        path_counts.reserve(est->paths.size());
        for (auto & path: est->paths) { path_counts.emplace_back(path.source_count); }

        planPathGroupPosteriors(&plans[k], est, est->paths.size(), path_counts, group_size, bounded, min_rel_likelihood);
    }

    auto * ctx = threadContext();

    // the dense matrices (rows x paths of every cluster, as denseMatrix makes them) are built on the device
    rpvg_b200_read_records records = arena.view();
    rpvg_b200_grouped * resident = nullptr;
    trace.lap("haplotypes: plans");

    if (rpvg_b200_grouped_raw_records(ctx, &records, path_offset.data(), &resident) != RPVG_B200_OK) { fatalBatch(ctx, "rpvg_b200_grouped_raw_records"); }

    trace.lap("haplotypes: upload + dense matrices");

    vector<PathClusterEstimates *> outs(n);
    for (size_t k = 0; k < n; ++k) { outs[k] = estimates[members[k]]; }

    runPosteriorRounds(ctx, resident, path_offset, &plans, outs);

    rpvg_b200_grouped_free(ctx, resident);
}

void PathPosteriorEstimator::estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs) {

    assert(estimates.size() == cluster_probs.size());

    for (auto & est: estimates) { est->resetEstimates(est->paths.size(), 1); }
    posteriorsBatch(estimates, cluster_probs, 1, false, 1);
}

double PathGroupPosteriorEstimator::clusterCost(const PathClusterEstimates & estimates, const vector<ReadPathProbabilities> & cluster_probs) const {

    const double paths = estimates.paths.size();
    return (group_size == 2 ? paths * (paths + 1) / 2 : paths) * max<double>(1, cluster_probs.size()) + PathEstimator::clusterCost(estimates, cluster_probs);
}

double NestedPathAbundanceEstimator::clusterCost(const PathClusterEstimates & estimates, const vector<ReadPathProbabilities> & cluster_probs) const {

    // at most one haplotype group per haplotype id of the cluster, at most as many as path combinations
    rpvg_b200_host::hash_set<uint32_t> ids;
    for (auto & path: estimates.paths) {
        for (auto & id: path.source_ids) { ids.insert(id); }
    }

    const double groups = max<double>(1, ids.size());
    return groups * (groups + 1) / 2 * max<double>(1, cluster_probs.size()) + PathEstimator::clusterCost(estimates, cluster_probs);
}

void PathGroupPosteriorEstimator::estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs) {

    assert(estimates.size() == cluster_probs.size());

    if (use_group_post_gibbs) {
        PathEstimator::estimateBatch(estimates, cluster_probs, rngs);
        return;
    }

    for (auto & est: estimates) { est->resetEstimates(0, 0); }

    // rpvg:src/path_posterior_estimator.cpp:5 — min_rel_likelihood of the bounded calculation
    posteriorsBatch(estimates, cluster_probs, group_size, group_size == 2, 1e-8);
}

// strains model (rpvg:src/path_abundance_estimator.cpp:217-295): covers of all clusters in one call, then the EM on every
// cluster's cover (partial matrix, normalise, collapse, EM, Gibbs) in one call
void MinimumPathAbundanceEstimator::estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs) {

    assert(estimates.size() == cluster_probs.size());
    assert(estimates.size() == rngs.size());

    vector<size_t> members;

    for (size_t i = 0; i < estimates.size(); ++i) {
        estimates[i]->resetEstimates(estimates[i]->paths.size(), 1);
        if (!cluster_probs[i]->empty()) { members.push_back(i); }
    }

    if (members.empty()) { return; }

    BatchArena arena;
    {
        vector<const vector<ReadPathProbabilities> *> member_probs;
        vector<uint32_t> member_paths;
        for (auto i: members) {
            member_probs.push_back(cluster_probs[i]);
            member_paths.push_back(estimates[i]->paths.size());
        }
        arena.build(threadContext(), member_probs, member_paths);
    }

    rpvg_b200_clusters clusters = arena.view();
    auto * ctx = threadContext();

    // rpvg:src/path_abundance_estimator.cpp:229-252
    vector<uint8_t> in_cover(arena.path_offset.back(), 0);
    if (rpvg_b200_path_cover(ctx, &clusters, in_cover.data()) != RPVG_B200_OK) { fatalBatch(ctx, "rpvg_b200_path_cover"); }

    // :254-282 one subset (the cover) per cluster that has one
    vector<uint32_t> subset_cluster, subset_paths, gibbs_counts;
    vector<uint64_t> subset_offset = {0}, gibbs_seeds;
    size_t total_values = 0;

    for (size_t k = 0; k < members.size(); ++k) {

        const size_t p0 = arena.path_offset[k], p1 = arena.path_offset[k + 1];
        const size_t before = subset_paths.size();

        for (size_t j = p0; j < p1; ++j) {
            if (in_cover[j]) { subset_paths.push_back(static_cast<uint32_t>(j - p0)); }
        }

        if (subset_paths.size() == before) { continue; }   // empty cover: the reset estimates stand

        subset_cluster.push_back(static_cast<uint32_t>(k));
        subset_offset.push_back(subset_paths.size());
        gibbs_counts.push_back(num_gibbs_samples);

        uint64_t seed = 0;
        if (num_gibbs_samples > 0) {
            mt19937 * mt_rng = rngs[members[k]];
            seed = (static_cast<uint64_t>((*mt_rng)()) << 32) | (*mt_rng)();
        }
        gibbs_seeds.push_back(seed);
        total_values += static_cast<size_t>(num_gibbs_samples) * (subset_paths.size() - before);
    }

    if (subset_cluster.empty()) { return; }

    rpvg_b200_subsets subsets = {static_cast<uint32_t>(subset_cluster.size()), subset_cluster.data(), subset_offset.data(), subset_paths.data()};

    rpvg_b200_em_params params;
    rpvg_b200_em_params_default(&params);
    params.max_em_its = max_em_its;
    params.max_rel_em_conv = max_rel_em_conv;

    vector<double> cover_abundances(subset_paths.size(), 0), noise_counts(subset_cluster.size(), 0), total_counts(subset_cluster.size(), 0);
    vector<double> gibbs_abundances(total_values), gibbs_noise(static_cast<size_t>(num_gibbs_samples) * subset_cluster.size());

    rpvg_b200_gibbs gibbs = {gibbs_counts.data(), gibbs_thin_its, gibbs_seeds.data(), gibbs_abundances.data(), gibbs_noise.data()};

    if (rpvg_b200_subset_em_gibbs(ctx, &clusters, &subsets, &params, prob_precision, num_gibbs_samples > 0 ? &gibbs : nullptr, cover_abundances.data(), noise_counts.data(),
                                  total_counts.data(), nullptr, nullptr) != RPVG_B200_OK) {
        fatalBatch(ctx, "rpvg_b200_subset_em");
    }

    size_t value_pos = 0;

    for (size_t s = 0; s < subset_cluster.size(); ++s) {

        PathClusterEstimates * est = estimates[members[subset_cluster[s]]];
        const size_t c0 = subset_offset[s], c1 = subset_offset[s + 1];

        if (num_gibbs_samples > 0) {

            CountSamples samples;
            samples.path_ids.assign(subset_paths.begin() + c0, subset_paths.begin() + c1);
            samples.noise_samples.assign(gibbs_noise.begin() + s * num_gibbs_samples, gibbs_noise.begin() + (s + 1) * num_gibbs_samples);
            samples.abundance_samples.assign(gibbs_abundances.begin() + value_pos, gibbs_abundances.begin() + value_pos + static_cast<size_t>(num_gibbs_samples) * (c1 - c0));
            value_pos += static_cast<size_t>(num_gibbs_samples) * (c1 - c0);

            est->gibbs_read_count_samples.emplace_back(std::move(samples));
        }

        // :286-292
        for (size_t i = c0; i < c1; ++i) { est->abundances.at(subset_paths[i]) += cover_abundances[i]; }

        est->noise_count = noise_counts[s];
        est->total_count = total_counts[s];
    }
}
