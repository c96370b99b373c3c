"""Haplotype probabilities and read counts per path (rpvg_b200_haplotype_marginals) against a restatement of the loop in
HaplotypeAbundanceEstimatesWriter::addEstimates (rpvg:src/threaded_output_writer.cpp:355-391). Bit-exact: the device adds
each path's contributions in the order of the sets, like the reference.
"""

import numpy as np
import pytest

from rpvg_b200 import capi

pytestmark = pytest.mark.gpu


def reference_marginals(path_offset, set_offset, member_offset, members, posteriors, abundances):
    """The reference's loop, set by set (rpvg:src/threaded_output_writer.cpp:361-385)."""
    probs = np.zeros(int(path_offset[-1]))
    counts = np.zeros(int(path_offset[-1]))
    for c in range(len(path_offset) - 1):
        p0 = int(path_offset[c])
        for i in range(int(set_offset[c]), int(set_offset[c + 1])):
            e0, e1 = int(member_offset[i]), int(member_offset[i + 1])
            for e in range(e0, e1):
                if e == e0 or members[e] != members[e - 1]:
                    probs[p0 + members[e]] += posteriors[i]
                counts[p0 + members[e]] += abundances[e]
    return probs, counts


def random_estimates(rng, n_clusters, ploidy, max_paths=40, max_sets=200, empty_every=7):
    path_offset, set_offset, member_offset = [0], [0], [0]
    members = []
    for c in range(n_clusters):
        n_paths = int(rng.integers(1, max_paths + 1))
        n_sets = 0 if (empty_every and c % empty_every == 3) else int(rng.integers(1, max_sets + 1))
        path_offset.append(path_offset[-1] + n_paths)
        for _ in range(n_sets):
            size = int(rng.integers(1, ploidy + 1))
            group = np.sort(rng.integers(0, n_paths, size=size))      # repeats (homozygous sets) end up adjacent
            members.extend(int(x) for x in group)
            member_offset.append(len(members))
        set_offset.append(len(member_offset) - 1)
    n_sets_total = len(member_offset) - 1
    posteriors = rng.uniform(0, 1, size=n_sets_total)
    posteriors[rng.uniform(size=n_sets_total) < 0.1] = 0.0
    abundances = rng.gamma(0.5, 50.0, size=len(members))
    abundances[rng.uniform(size=len(members)) < 0.2] = 0.0
    return (np.array(path_offset, dtype=np.uint64), np.array(set_offset, dtype=np.uint64), np.array(member_offset, dtype=np.uint64),
            np.array(members, dtype=np.uint32), posteriors, abundances)


@pytest.mark.parametrize("ploidy", [1, 2, 3])
def test_marginals_match_reference_loop(ctx, ploidy):
    rng = np.random.default_rng(100 + ploidy)
    args = random_estimates(rng, n_clusters=300, ploidy=ploidy)
    probs, counts = ctx.haplotype_marginals(*args)
    ref_probs, ref_counts = reference_marginals(*args)
    assert np.array_equal(probs, ref_probs)
    assert np.array_equal(counts, ref_counts)


def test_marginals_many_sets_on_one_path(ctx):
    """Long accumulation chains (thousands of sets on few paths): the order of the additions is what makes it bit-exact."""
    rng = np.random.default_rng(7)
    args = random_estimates(rng, n_clusters=5, ploidy=2, max_paths=3, max_sets=5000, empty_every=0)
    probs, counts = ctx.haplotype_marginals(*args)
    ref_probs, ref_counts = reference_marginals(*args)
    assert np.array_equal(probs, ref_probs)
    assert np.array_equal(counts, ref_counts)


def test_marginals_edge_cases(ctx):
    # no sets at all: zeros for every path
    probs, counts = ctx.haplotype_marginals([0, 3, 5], [0, 0, 0], [0], [], [], [])
    assert np.array_equal(probs, np.zeros(5)) and np.array_equal(counts, np.zeros(5))

    # a homozygous set counts its posterior once and both abundances; member order inside a set is kept
    probs, counts = ctx.haplotype_marginals([0, 2], [0, 2], [0, 2, 4], [1, 1, 0, 1], [0.25, 0.5], [3.0, 4.0, 1.5, 2.5])
    assert np.array_equal(probs, np.array([0.5, 0.75]))
    assert np.array_equal(counts, np.array([1.5, 9.5]))

    # a member outside the cluster's paths is an error, and the context stays usable
    with pytest.raises(capi.RpvgB200Error):
        ctx.haplotype_marginals([0, 2], [0, 1], [0, 2], [0, 2], [1.0], [1.0, 1.0])
    probs, _ = ctx.haplotype_marginals([0, 2], [0, 1], [0, 2], [0, 1], [1.0], [1.0, 1.0])
    assert np.array_equal(probs, np.array([1.0, 1.0]))


def test_marginals_of_nested_estimates(ctx):
    """Estimates of the nested model on synthetic clusters: per cluster the read counts add up to the abundances and the
    haplotype probabilities of the paths of one transcript add up to at most ploidy."""
    from rpvg_b200 import synth
    from rpvg_b200.arena import concat
    from oracle import loader
    import os

    rng = np.random.default_rng(5)
    arena = concat([synth.haplotype_cluster(rng, n_transcripts=2, paths_per_transcript=(2, 6), n_haplotypes=12, n_reads=300) for _ in range(6)])
    b200 = loader.load_harness(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "libharness_b200.so"))
    _, results = loader.harness_estimate_batch(b200, loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2), arena)

    path_offset, set_offset, member_offset, members, posteriors, abundances = [0], [0], [0], [], [], []
    for c, res in enumerate(results):
        n_paths = int(arena.path_offset[c + 1] - arena.path_offset[c])
        path_offset.append(path_offset[-1] + n_paths)
        for group, posterior in zip(res["sets"], res["posteriors"]):
            members.extend(int(x) for x in group)
            member_offset.append(len(members))
            posteriors.append(float(posterior))
        abundances.extend(float(x) for x in res["abundances"])
        set_offset.append(len(member_offset) - 1)

    probs, counts = ctx.haplotype_marginals(path_offset, set_offset, member_offset, members, posteriors, abundances)
    ref_probs, ref_counts = reference_marginals(path_offset, set_offset, member_offset, np.array(members), posteriors, abundances)
    assert np.array_equal(probs, ref_probs) and np.array_equal(counts, ref_counts)
    for c, res in enumerate(results):
        assert np.isclose(counts[path_offset[c]:path_offset[c + 1]].sum(), np.sum(res["abundances"]), rtol=1e-12)
