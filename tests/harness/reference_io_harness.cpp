// Reference-only oracles (test infrastructure, not product). See reference_io_harness.h.
// Compiled with -fno-access-control against rpvg's own, unmodified sources.

#include <cstdint>
#include <string>
#include <vector>

#include "sparsepp/spp.h"

#include "alignment_path.hpp"
#include "fragment_length_dist.hpp"
#include "path_cluster_estimates.hpp"
#include "read_path_probabilities.hpp"
#include "threaded_output_writer.hpp"

#include "reference_io_harness.h"

using namespace std;

extern "C" {

double ref_frag_log_prob(double mean, double sd, uint32_t sd_max_multi, uint32_t value) {

    FragmentLengthDist dist(mean, sd, sd_max_multi);
    return dist.logProb(value);
}

uint64_t ref_add_path_probs(const HarnessAlignments * al, double frag_mean, double frag_sd, uint32_t sd_max_multi, int32_t is_single_end, double min_noise_prob,
                            double prob_precision, int32_t collapse_groups, double * out_noise, uint64_t * out_group_ptr, double * out_group_prob,
                            uint64_t * out_group_id_ptr, uint32_t * out_group_ids) {

    FragmentLengthDist fragment_length_dist(frag_mean, frag_sd, sd_max_multi);

    spp::sparse_hash_map<uint32_t, uint32_t> clustered_path_index;
    vector<PathInfo> cluster_paths;
    spp::sparse_hash_map<string, uint32_t> group_name_index;

    for (uint32_t p = 0; p < al->num_paths; ++p) {

        clustered_path_index.emplace(p, p);

        const string name = collapse_groups ? "g" + to_string(al->path_group[p]) : "p" + to_string(p);
        cluster_paths.emplace_back(PathInfo(name));
        cluster_paths.back().effective_length = al->effective_length[p];
        cluster_paths.back().source_count = al->source_count ? al->source_count[p] : 1;

        if (collapse_groups) { group_name_index.emplace(name, al->path_group[p]); }
    }

    uint64_t n_groups = 0, n_ids = 0;
    out_group_ptr[0] = 0;
    out_group_id_ptr[0] = 0;

    for (uint32_t r = 0; r < al->num_reads; ++r) {

        vector<AlignmentPath> align_paths;
        vector<vector<gbwt::size_type> > align_paths_ids;

        for (uint64_t a = al->align_ptr[r]; a < al->align_ptr[r + 1]; ++a) {

            align_paths.emplace_back(make_pair(gbwt::SearchState(), 0), true, al->min_mapq[r], al->score_sum[a], static_cast<uint16_t>(al->align_length[a]),
                                     static_cast<uint16_t>(al->frag_length[a]));
            align_paths_ids.emplace_back(al->ids + al->id_ptr[a], al->ids + al->id_ptr[a + 1]);
        }

        ReadPathProbabilities probs(1, prob_precision);
        probs.addPathProbs(align_paths, align_paths_ids, clustered_path_index, cluster_paths, fragment_length_dist, is_single_end != 0, min_noise_prob, collapse_groups != 0,
                           group_name_index);

        out_noise[r] = probs.noiseProb();

        for (auto & group: probs.pathProbs()) {
            out_group_prob[n_groups] = group.first;
            for (auto id: group.second) { out_group_ids[n_ids++] = id; }
            out_group_id_ptr[++n_groups] = n_ids;
        }
        out_group_ptr[r + 1] = n_groups;
    }

    return n_groups;
}

int32_t ref_write_probs(const char * prefix, double prob_precision, uint32_t n, const HarnessCluster * clusters, const uint32_t * const * length,
                        const double * const * effective_length) {

    ProbabilityClusterWriter writer(prefix, 1, prob_precision);

    for (uint32_t c = 0; c < n; ++c) {

        const HarnessCluster & cl = clusters[c];

        vector<PathInfo> paths;
        for (uint32_t p = 0; p < cl.num_paths; ++p) {
            paths.emplace_back(PathInfo("c" + to_string(c) + "_p" + to_string(p)));
            paths.back().length = length[c][p];
            paths.back().effective_length = effective_length[c][p];
        }

        vector<ReadPathProbabilities> cluster_probs;
        for (uint32_t r = 0; r < cl.num_reads; ++r) {

            cluster_probs.emplace_back(ReadPathProbabilities(cl.read_count[r], prob_precision));
            auto & probs = cluster_probs.back();
            probs.noise_prob = cl.noise_prob[r];

            for (uint64_t e = cl.entry_ptr[r]; e < cl.entry_ptr[r + 1]; ++e) {
                probs.path_probs.emplace_back(cl.entry_prob[e], vector<uint32_t>(cl.entry_path_ids + cl.entry_path_ptr[e], cl.entry_path_ids + cl.entry_path_ptr[e + 1]));
            }
        }

        writer.addCluster(cluster_probs, paths);
    }

    writer.close();
    return 0;
}

int32_t ref_write_estimates(int32_t kind, const char * prefix, uint32_t ploidy, double min_posterior, double total_transcript_count, uint32_t n,
                            const HarnessEstimates * estimates, uint32_t unaligned_read_count) {

    vector<pair<uint32_t, PathClusterEstimates> > all;
    all.reserve(n);

    for (uint32_t c = 0; c < n; ++c) {

        const HarnessEstimates & he = estimates[c];
        all.emplace_back(he.cluster_id, PathClusterEstimates());
        auto & est = all.back().second;

        for (uint32_t p = 0; p < he.num_paths; ++p) {
            est.paths.emplace_back(PathInfo("c" + to_string(he.cluster_id) + "_p" + to_string(p)));
            est.paths.back().length = he.length[p];
            est.paths.back().effective_length = he.effective_length[p];
        }

        for (uint64_t i = 0; i < he.n_sets; ++i) {
            est.path_group_sets.emplace_back(he.set_ids + he.set_ptr[i], he.set_ids + he.set_ptr[i + 1]);
        }

        est.posteriors.assign(he.posteriors, he.posteriors + he.n_sets);
        est.abundances.assign(he.abundances, he.abundances + he.n_abundances);
        est.noise_count = he.noise_count;
        est.total_count = he.total_count;
    }

    if (kind == 0) {

        AbundanceEstimatesWriter writer(prefix, 1, total_transcript_count);
        writer.addEstimates(all);
        writer.addNoiseTranscript(unaligned_read_count);
        writer.close();

    } else if (kind == 1) {

        HaplotypeAbundanceEstimatesWriter writer(prefix, 1, ploidy, total_transcript_count);
        writer.addEstimates(all);
        writer.addNoiseTranscript(unaligned_read_count);
        writer.close();

    } else if (kind == 2) {

        JointHaplotypeAbundanceEstimatesWriter writer(prefix, 1, ploidy, min_posterior, total_transcript_count);
        writer.addEstimates(all);
        writer.addNoiseTranscript(unaligned_read_count);
        writer.close();

    } else {

        JointHaplotypeEstimatesWriter writer(prefix, 1, ploidy, min_posterior);
        writer.addEstimates(all);
        writer.close();
    }

    return 0;
}

}
