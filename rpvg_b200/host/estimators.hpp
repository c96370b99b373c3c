// rpvg_b200 host side: drop-in estimator classes with the reference's names, constructor and estimate() signatures
// (rpvg:src/path_estimator.hpp:16-49, path_abundance_estimator.hpp:18-79, path_posterior_estimator.hpp:18-42), so that
// rpvg:src/main.cpp:766-788,977 compiles unchanged. Everything O(data) runs on the GPU through the C ABI
// (include/rpvg_b200.h); what stays here is the O(#sets) combinatorics whose order is observable in the output.
//
// Inside rpvg's tree this header takes the place of path_estimator.hpp, path_abundance_estimator.hpp and
// path_posterior_estimator.hpp; compile with -DRPVG_B200_WITH_RPVG_HEADERS: the value types (PathInfo, PathClusterEstimates,
// ReadPathProbabilities, the Eigen typedefs of utils.hpp) then come from rpvg's own headers. The test infrastructure builds
// that configuration against the reference tree (the reference's own Catch2 test of the path cover compiled against these
// classes). Stand-alone (this repository's tests) the same public types are declared below.
#ifndef RPVG_B200_HOST_ESTIMATORS_HPP
#define RPVG_B200_HOST_ESTIMATORS_HPP

#include <cstdint>
#include <random>
#include <string>
#include <utility>
#include <vector>

#ifdef RPVG_B200_WITH_RPVG_HEADERS

#include "sparsepp/spp.h"
#include "path_cluster_estimates.hpp"
#include "read_path_probabilities.hpp"
#include "utils.hpp"

#define RPVG_B200_HAVE_SPARSEPP 1

namespace rpvg_b200_host {
template <class K> using hash_set = spp::sparse_hash_set<K>;
template <class K, class V> using hash_map = spp::sparse_hash_map<K, V>;
}

using namespace std;

#else

#if __has_include("sparsepp/spp.h")
#include "sparsepp/spp.h"
#define RPVG_B200_HAVE_SPARSEPP 1
namespace rpvg_b200_host {
template <class K> using hash_set = spp::sparse_hash_set<K>;
template <class K, class V> using hash_map = spp::sparse_hash_map<K, V>;
}
#else
#include <unordered_map>
#include <unordered_set>
namespace rpvg_b200_host {
template <class K> using hash_set = std::unordered_set<K>;
template <class K, class V> using hash_map = std::unordered_map<K, V>;
}
#endif

using namespace std;

// rpvg:src/path_cluster_estimates.hpp:14-32
struct PathInfo {

    string name;
    uint32_t group_id;

    uint32_t source_count;
    rpvg_b200_host::hash_set<uint32_t> source_ids;

    uint32_t length;
    double effective_length;

    PathInfo(const string & name_in) : name(name_in), group_id(0), source_count(1), length(0), effective_length(0) {}
};

// rpvg:src/path_cluster_estimates.hpp:34-42
struct CountSamples {

    vector<uint32_t> path_ids;

    vector<double> noise_samples;
    vector<double> abundance_samples;
};

// rpvg:src/path_cluster_estimates.hpp:44-110
struct PathClusterEstimates {

    vector<PathInfo> paths;

    vector<vector<uint32_t> > path_group_sets;

    vector<double> posteriors;
    vector<double> abundances;

    double noise_count;
    double total_count;

    vector<CountSamples> gibbs_read_count_samples;

    PathClusterEstimates() : noise_count(0), total_count(0) {}

    // all size-k multisets of {0..n-1} in lexicographic order; posteriors and abundances zeroed
    void resetEstimates(uint32_t num_components, const uint32_t group_size);
};

// rpvg:src/read_path_probabilities.hpp:19-44 (the record; addPathProbs stays in rpvg)
class ReadPathProbabilities {

    public:

        ReadPathProbabilities() : read_count(0), noise_prob(1), prob_precision(0) {}
        ReadPathProbabilities(const uint32_t read_count_in, const double prob_precision_in) : read_count(read_count_in), noise_prob(1), prob_precision(prob_precision_in) {}

        uint32_t readCount() const { return read_count; }
        double noiseProb() const { return noise_prob; }
        const vector<pair<double, vector<uint32_t> > > & pathProbs() const { return path_probs; }

        void addReadCount(const uint32_t read_count_in) { read_count += read_count_in; }

    private:

        uint32_t read_count;
        double noise_prob;
        vector<pair<double, vector<uint32_t> > > path_probs;

        double prob_precision;
};

#endif

// rpvg:src/path_estimator.hpp:51-68 — the hash the nested model's maps of path lists are keyed with; it decides their
// iteration order, which is observable in the output.
namespace std {

    template<>
    struct hash<vector<uint32_t> > {

        size_t operator()(vector<uint32_t> const & vec) const {

            size_t seed = 0;
#ifdef RPVG_B200_HAVE_SPARSEPP
            for (auto & val: vec) { spp::hash_combine(seed, val); }
#else
            for (auto & val: vec) { seed ^= std::hash<uint32_t>()(val) + 0x9e3779b9 + (seed << 6) + (seed >> 2); }
#endif
            return seed;
        }
    };
}

struct rpvg_b200_ctx;
struct rpvg_b200_grouped;

class PathEstimator {

    public:

        PathEstimator(const double prob_precision_in);
        virtual ~PathEstimator();

        // The reference's interface (rpvg:src/path_estimator.hpp:23), called once per cluster from rpvg's OpenMP loop on one
        // shared object (rpvg:src/main.cpp:829-830, 977). A device call per cluster costs a few hundred microseconds of
        // launches, copies and synchronisation whatever the cluster's size, so concurrent callers are combined: while one
        // thread is on the device, the clusters the other threads bring queue up, and whoever finds the device free next
        // takes all of them through estimateBatch in one go (each cluster with its own generator; results identical to
        // one-by-one calls). A thread that calls alone gets estimateOne directly. RPVG_B200_COMBINE=0 switches it off.
        virtual void estimate(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng);

        // whether estimate() combines concurrent callers (measured per class on 4000 small clusters, 16 threads: it halves the
        // time of the abundance classes and doubles that of the posterior classes, whose single call is short)
        virtual bool combinesCallers() const { return true; }

        // one cluster on the calling thread's context: what the reference's estimate() of the class does
        virtual void estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) = 0;

        // Not in rpvg: the cluster driver loop (rpvg:src/main.cpp:829-977) re-stated as one call for many clusters, so that
        // the device sees a few large launches instead of several small ones per cluster. Same
        // results, cluster by cluster, as estimate(estimates[i], *cluster_probs[i], rngs[i]); the default does just that.
        virtual void estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs);

        // Not in rpvg: estimateBatch over several GPUs of the box. Clusters are independent: they are dealt to
        // the devices by longest-processing-time-first on a per-cluster cost (clusterCost), one host thread with its own
        // context and streams per device runs estimateBatch on its share, and every result lands in the caller's own
        // PathClusterEstimates object — the original order, nothing to gather, no collective. The results are those of
        // the single-device call, bit for bit (a cluster's computation does not depend on what else is in its batch).
        // num_devices <= 0: $RPVG_B200_DEVICES ("all" or a count), else 1; never more than the visible devices.
        void estimateBatchOnDevices(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs, int num_devices = 0);

    private:

        // callers waiting for the device (estimate)
        struct Combiner;
        Combiner * combiner;

    protected:

        // what a cluster costs, for the split over devices: entries x expected EM iterations by default
        virtual double clusterCost(const PathClusterEstimates & estimates, const vector<ReadPathProbabilities> & cluster_probs) const;

        // device of the calling thread's context: $RPVG_B200_DEVICE (default 0) unless estimateBatchOnDevices set it
        static int & threadDevice();

        const double prob_precision;

        // The posterior calculations below in stages, so that the device requests of many clusters can be put into one
        // call: a plan holds the path sets whose log-likelihoods are wanted next (flat_sets, set_size paths each).
        struct GroupPosteriorPlan {

            uint32_t num_paths = 0;
            uint32_t group_size = 0;
            bool bounded = false;
            double min_log_likelihood_diff = 0;
            vector<double> path_log_freqs;

            bool device_pairs = false;   // bounded: the pair round is left to rpvg_b200_grouped_bounded_pairs (batch.cpp)
            int stage = 0;
            uint32_t set_size = 0;
            vector<uint32_t> flat_sets;

            vector<pair<double, uint32_t> > marginal_posteriors;
            PathClusterEstimates marginal_estimates;
        };

        void planPathGroupPosteriors(GroupPosteriorPlan * plan, PathClusterEstimates * path_cluster_estimates, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size, const bool bounded, const double min_rel_likelihood) const;

        // Consumes the log-likelihoods of plan->flat_sets; true = the plan holds another request, false = the posteriors
        // are in path_cluster_estimates.
        bool advancePathGroupPosteriors(GroupPosteriorPlan * plan, PathClusterEstimates * path_cluster_estimates, const double * loglik) const;

        // Last stage of a bounded plan with device_pairs: the reference's walk over the candidate pairs the device kept
        // (index in the (i, j >= i) enumeration of plan->marginal_posteriors, log-likelihood sum), in pair order.
        void finishBoundedPathGroupPosteriors(GroupPosteriorPlan * plan, PathClusterEstimates * path_cluster_estimates, const size_t n_candidates, const uint32_t * cand_pair, const double * cand_loglik) const;

        // Full (any k) or Bounded (k = 2) posteriors over the raw probability matrices of many clusters: the matrices go
        // to the device once, every round of log-likelihoods is one call for all clusters (batch.cpp).
        // The device rounds of a batch of plans (all of the same kind) on resident matrices: one log-likelihood call for
        // the first request of all plans; for bounded plans the pair walk on the device, then the exact walk over its
        // candidates. group_offset [plans + 1]: columns (paths or groups) of each plan's matrix.
        void runPosteriorRounds(rpvg_b200_ctx * ctx, rpvg_b200_grouped * resident, const vector<uint64_t> & group_offset, vector<GroupPosteriorPlan> * plans, const vector<PathClusterEstimates *> & outs) const;

        void posteriorsBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const uint32_t group_size, const bool bounded, const double min_rel_likelihood) const;

        // The context device calls go to: the combiner's while a thread works for it, else one per calling thread
        // (estimateBatch and estimateOne are re-entrant on a shared object like the reference's estimate, rpvg:src/main.cpp:829-830).
        static rpvg_b200_ctx * threadContext();
        static rpvg_b200_ctx *& contextOverride();

        // Dense column-major matrix as constructProbabilityMatrix leaves it (rpvg:src/path_estimator.cpp:55-77).
        static void denseMatrix(vector<double> * matrix, vector<double> * noise, vector<double> * counts, const vector<ReadPathProbabilities> & cluster_probs, const uint32_t num_paths);

        // rpvg:src/path_estimator.cpp:332-377 / 379-473 with the log-likelihoods computed on the device.
        void calculatePathGroupPosteriorsFull(PathClusterEstimates * path_cluster_estimates, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size) const;
        void calculatePathGroupPosteriorsBounded(PathClusterEstimates * path_cluster_estimates, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size, const double min_rel_likelihood) const;

        // rpvg:src/path_estimator.cpp:475-589. The chain runs on the host with the caller's mt19937 and libstdc++'s
        // discrete_distribution, in the reference's consumption order (the RNG contract of rpvg:src/main.cpp:976); each
        // conditional distribution it needs (N log-likelihoods over all read classes) is computed on the device.
        void estimatePathGroupPosteriorsGibbs(PathClusterEstimates * path_cluster_estimates, const vector<double> & matrix, const vector<double> & noise, const vector<double> & counts, const uint32_t num_paths, const vector<uint32_t> & path_counts, const uint32_t group_size, mt19937 * mt_rng) const;
};

class PathAbundanceEstimator : public PathEstimator {

    public:

        PathAbundanceEstimator(const uint32_t max_em_its_in, const double max_rel_em_conv_in, const uint32_t num_gibbs_samples_in, const uint32_t gibbs_thin_its_in, const double prob_precision);
        virtual ~PathAbundanceEstimator() {};

        void estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng);

        // all clusters in one device batch (EM and, if asked for, the Gibbs read-count samples)
        void estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs);

    protected:

        const uint32_t max_em_its;
        const double max_rel_em_conv;

        const uint32_t num_gibbs_samples;
        const uint32_t gibbs_thin_its;
};

class MinimumPathAbundanceEstimator : public PathAbundanceEstimator {

    public:

        MinimumPathAbundanceEstimator(const uint32_t max_em_its, const double max_rel_em_conv, const uint32_t num_gibbs_samples, const uint32_t gibbs_thin_its, const double prob_precision);
        ~MinimumPathAbundanceEstimator() {};

        void estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng);

#ifdef RPVG_B200_WITH_RPVG_HEADERS
        // rpvg:src/path_abundance_estimator.hpp:48 — public in the reference (its test calls it); the picks run on the device
        vector<uint32_t> weightedMinimumPathCover(const Utils::ColMatrixXb & read_path_cover, const Utils::RowVectorXd & read_counts, const Utils::RowVectorXd & path_weights) const;
#endif
        // the same on flat arrays: cover column-major, rows x n_paths, non-zero = the read class has the path
        vector<uint32_t> weightedMinimumPathCover(const uint32_t rows, const uint32_t n_paths, const uint8_t * read_path_cover, const double * read_counts, const double * path_weights) const;

        // two device calls for the whole batch: the path covers of all clusters, then the EMs (and Gibbs samples) on them
        void estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs);
};

class NestedPathAbundanceEstimator : public PathAbundanceEstimator {

    public:

        NestedPathAbundanceEstimator(const uint32_t group_size_in, const double min_hap_prob_in, const bool infer_collapsed_in, const bool use_group_post_gibbs_in, const uint32_t max_em_its, const double max_rel_em_conv, const uint32_t num_gibbs_samples, const uint32_t gibbs_thin_its, const double prob_precision);
        ~NestedPathAbundanceEstimator() {};

        void estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng);

        // Collapsed mode without --use-hap-gibbs: four device calls for the whole batch (grouped + collapsed matrices,
        // one or two rounds of log-likelihoods, all sub-EMs of all clusters) instead of four per cluster.
        void estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs);

    protected:

        // haplotype combinations x read classes dominate this model
        double clusterCost(const PathClusterEstimates & estimates, const vector<ReadPathProbabilities> & cluster_probs) const;

    private:

        const uint32_t group_size;
        const double min_hap_prob;

        const bool infer_collapsed;
        const bool use_group_post_gibbs;

        typedef rpvg_b200_host::hash_map<vector<uint32_t>, double> SubsetSamples;

        void inferAbundancesIndependentGroups(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) const;
        void inferAbundancesCollapsedGroups(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng) const;

        // group posteriors on the grouped + normalised + collapsed matrix of the given path groups
        void groupPosteriors(PathClusterEstimates * group_estimates, const vector<ReadPathProbabilities> & cluster_probs, const uint32_t num_paths, const vector<vector<uint32_t> > & path_groups, const vector<uint32_t> & group_counts, mt19937 * mt_rng) const;

        vector<vector<uint32_t> > findPathGroups(const vector<PathInfo> & paths) const;
        pair<vector<vector<uint32_t> >, vector<uint32_t> > findPathSourceGroups(const vector<PathInfo> & paths) const;

        void inferPathSubsetAbundance(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng, const SubsetSamples & path_subset_samples) const;

        // inferPathSubsetAbundance in two halves around its device call (the sub-EMs of the subsets)
        struct SubsetJob {
            vector<uint32_t> paths;                // sorted, duplicates kept
            double prob;
            vector<uint32_t> collapsed;            // duplicates removed
            vector<uint32_t> multiplicity;         // per collapsed path
            uint32_t gibbs_samples;                // this subset's share of --num-gibbs-samples
            uint64_t gibbs_seed;
        };

        vector<SubsetJob> planPathSubsets(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng, const SubsetSamples & path_subset_samples) const;
        void finishPathSubsets(PathClusterEstimates * path_cluster_estimates, const vector<SubsetJob> & jobs, const double * abundances, const double * noise_counts, const double * gibbs_abundances, const double * gibbs_noise) const;

        // selectPathSubsetIndices of the collapsed mode (rpvg:src/path_abundance_estimator.cpp:569-606)
        SubsetSamples collapsedSubsetSamples(const PathClusterEstimates & group_estimates, const vector<vector<uint32_t> > & path_groups) const;
};

class PathPosteriorEstimator : public PathEstimator {

    public:

        PathPosteriorEstimator(const double prob_precision);
        virtual ~PathPosteriorEstimator() {};

        void estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng);
        void estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs);

        bool combinesCallers() const { return false; }
};

class PathGroupPosteriorEstimator : public PathPosteriorEstimator {

    public:

        PathGroupPosteriorEstimator(const uint32_t group_size_in, const bool use_group_post_gibbs_in, const double prob_precision);
        ~PathGroupPosteriorEstimator() {};

        void estimateOne(PathClusterEstimates * path_cluster_estimates, const vector<ReadPathProbabilities> & cluster_probs, mt19937 * mt_rng);

        // (--use-hap-gibbs samples on the host between the device calls of a cluster: looped)
        void estimateBatch(const vector<PathClusterEstimates *> & estimates, const vector<const vector<ReadPathProbabilities> *> & cluster_probs, const vector<mt19937 *> & rngs);

    protected:

        // path combinations x read classes
        double clusterCost(const PathClusterEstimates & estimates, const vector<ReadPathProbabilities> & cluster_probs) const;

    private:

        const uint32_t group_size;
        const bool use_group_post_gibbs;
};

#endif
