"""The drop-in estimator classes (rpvg_b200/host, same names and signatures as the reference's) against the reference's
own classes, through one harness source compiled against both (tests/harness/estimator_harness.cpp), and against the
frozen reference output in tests/golden.

Tolerances (north_star): read counts 1e-6 relative with the identical set of paths above the threshold; haplotype
posteriors 1e-9 absolute; set lists identical in content and order.
"""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth
from rpvg_b200.arena import concat
from tests.golden_util import cluster_results, load

pytestmark = pytest.mark.gpu

POST_TOL = 1e-9
REL_TOL = 1e-6


def assert_same_posteriors(mine, ref_sets, ref_post):
    assert mine["sets"] == ref_sets
    assert len(mine["posteriors"]) == len(ref_post)
    if len(ref_post):
        assert np.max(np.abs(mine["posteriors"] - ref_post)) < POST_TOL


def assert_same_abundances(mine, ref_ab, ref_noise, ref_total):
    assert np.array_equal(mine["abundances"] == 0, ref_ab == 0)
    nz = ref_ab != 0
    if nz.any():
        assert np.max(np.abs(mine["abundances"][nz] - ref_ab[nz]) / ref_ab[nz]) < REL_TOL
    assert abs(mine["noise_count"] - ref_noise) <= REL_TOL * max(1.0, ref_total)
    assert mine["total_count"] == ref_total


def hap_arena(seed, n=8):
    rng = np.random.default_rng(seed)
    return concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 4)), paths_per_transcript=(2, 8), n_haplotypes=30,
                                           n_reads=int(rng.integers(80, 400))) for _ in range(n)])


def test_transcripts_class_matches_golden(b200_harness):
    arena, z = load("transcripts_em.npz")
    _, results = loader.harness_estimate_batch(b200_harness, loader.make_config(loader.MODEL_TRANSCRIPTS), arena, num_threads=4)
    for c, res in enumerate(results):
        sets, post, ab, noise, total = cluster_results(z, "", c)
        assert res["sets"] == sets and np.all(res["posteriors"] == 0)
        assert_same_abundances(res, ab, noise, total)


@pytest.mark.parametrize("prefix,model,k", [("path_posterior_", loader.MODEL_PATH_POSTERIOR, 1), ("haplotypes_", loader.MODEL_HAPLOTYPES, 2)])
def test_posterior_classes_match_golden(b200_harness, prefix, model, k):
    arena, z = load("haplotype_models.npz")
    _, results = loader.harness_estimate_batch(b200_harness, loader.make_config(model, group_size=k), arena, rng_seed=7, num_threads=2)
    for c, res in enumerate(results):
        sets, post, ab, _, _ = cluster_results(z, prefix, c)
        assert_same_posteriors(res, sets, post)
        assert len(res["abundances"]) == len(ab)


@pytest.mark.parametrize("k", [1, 2, 3])
def test_haplotypes_class_matches_reference_class(b200_harness, reference, k):
    """PathGroupPosteriorEstimator: ploidy 1 and 3 -> Full over all multisets, ploidy 2 -> Bounded (min_rel_likelihood 1e-8)."""
    arena = hap_arena(300 + k)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=k)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena)
    for a, b in zip(mine, ref):
        assert_same_posteriors(a, b["sets"], b["posteriors"])
        assert abs(a["posteriors"].sum() - 1) < 1e-9


def test_empty_cluster_and_single_path(b200_harness, reference):
    arena = hap_arena(5, n=2)
    empty = arena.select([0])
    empty.read_offset[:] = 0
    empty.row_ptr = empty.row_ptr[:1]
    for model in (loader.MODEL_TRANSCRIPTS, loader.MODEL_HAPLOTYPES, loader.MODEL_PATH_POSTERIOR):
        cfg = loader.make_config(model)
        a = loader.harness_estimate(b200_harness, cfg, empty, 0)
        b = loader.harness_estimate(reference, cfg, empty, 0)
        assert a["sets"] == b["sets"] and np.array_equal(a["posteriors"], b["posteriors"]) and np.array_equal(a["abundances"], b["abundances"])


@pytest.mark.parametrize("k", [2, 4])
def test_haplotypes_gibbs_replays_the_reference_chain(b200_harness, reference, k):
    """--use-hap-gibbs: the chain runs on the host with the caller's mt19937 (the reference's RNG contract); its
    conditionals come from the device. Same seed -> same visited sets in the same order, same visit frequencies."""
    arena = hap_arena(700 + k, n=4)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=k, use_group_post_gibbs=True)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=5, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=5)
    agree = 0
    for a, b in zip(mine, ref):
        assert abs(a["posteriors"].sum() - 1) < 1e-9
        if a["sets"] == b["sets"] and np.max(np.abs(a["posteriors"] - b["posteriors"])) < 1e-12:
            agree += 1
        else:   # a discrete draw flipped on a last-bit difference: compare in distribution
            da = dict(zip(a["sets"], a["posteriors"])); db = dict(zip(b["sets"], b["posteriors"]))
            assert sum(abs(da.get(s, 0) - db.get(s, 0)) for s in set(da) | set(db)) < 0.2
    assert agree >= len(ref) - 1
