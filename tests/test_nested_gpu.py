"""Haplotype-transcripts model (NestedPathAbundanceEstimator) on the GPU against the reference's own class, through the
same harness source, and against the frozen reference output. The output tuples must come in the reference's order (it
is the iteration order of its hash maps); posteriors within 1e-9 absolute, read counts within 1e-6 relative."""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth
from rpvg_b200.arena import concat
from tests.golden_util import cluster_results, load

pytestmark = pytest.mark.gpu


def compare(mine, ref_sets, ref_post, ref_ab, ref_noise, ref_total):
    assert mine["sets"] == ref_sets, "path tuples / their order differ"
    assert np.max(np.abs(mine["posteriors"] - ref_post), initial=0) < 1e-9
    assert len(mine["abundances"]) == len(ref_ab)
    scale = max(1.0, ref_total)
    # abundances are posterior-weighted sums over sub-EMs: compare relative to the cluster's read count
    assert np.max(np.abs(mine["abundances"] - ref_ab), initial=0) <= 1e-6 * scale
    big = ref_ab > 1e-3 * scale
    if big.any():
        assert np.max(np.abs(mine["abundances"][big] - ref_ab[big]) / ref_ab[big]) < 1e-6
    assert abs(mine["noise_count"] - ref_noise) <= 1e-6 * scale
    assert mine["total_count"] == ref_total


def hap_arena(seed, n=6):
    rng = np.random.default_rng(seed)
    return concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 4)), paths_per_transcript=(2, 8), n_haplotypes=30,
                                           n_reads=int(rng.integers(100, 500))) for _ in range(n)])


def test_nested_collapsed_matches_golden(b200_harness):
    arena, z = load("haplotype_models.npz")
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
    _, results = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=7, num_threads=2)
    for c, res in enumerate(results):
        compare(res, *cluster_results(z, "nested_", c))


@pytest.mark.parametrize("k", [1, 2, 3])
def test_nested_collapsed_matches_reference_class(b200_harness, reference, k):
    arena = hap_arena(40 + k)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=k)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena)
    for a, b in zip(mine, ref):
        compare(a, b["sets"], b["posteriors"], b["abundances"], b["noise_count"], b["total_count"])
        assert abs(a["abundances"].sum() + a["noise_count"] - a["total_count"]) <= 1e-6 * a["total_count"]


def test_nested_independent_groups_matches_reference_class(b200_harness, reference):
    """--ind-hap-inference: per-transcript posteriors, joint samples drawn with the caller's mt19937 on the host."""
    arena = hap_arena(50, n=4)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2, infer_collapsed=False, min_hap_prob=0.01)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=11, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=11)
    for a, b in zip(mine, ref):
        compare(a, b["sets"], b["posteriors"], b["abundances"], b["noise_count"], b["total_count"])
