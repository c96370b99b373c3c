"""estimateBatch of the drop-in classes (the cluster driver loop as one call): the results must be
those of estimate() cluster by cluster — same tuples in the same order, same numbers — and therefore the reference's."""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import capi, synth
from rpvg_b200.arena import concat

pytestmark = pytest.mark.gpu


def hap_arena(seed, n):
    rng = np.random.default_rng(seed)
    return concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 4)), paths_per_transcript=(2, 8), n_haplotypes=30,
                                           n_reads=int(rng.integers(100, 500))) for _ in range(n)])


def same(a, b, exact=True):
    assert a["sets"] == b["sets"]
    for key in ("posteriors", "abundances"):
        if exact:
            assert np.array_equal(a[key], b[key]), key
        else:
            assert np.allclose(a[key], b[key], rtol=1e-9, atol=1e-12), key
    assert a["total_count"] == b["total_count"]
    assert a["noise_count"] == b["noise_count"] if exact else abs(a["noise_count"] - b["noise_count"]) <= 1e-9 * max(1.0, b["total_count"])


@pytest.mark.parametrize("k, gibbs", [(2, 0), (1, 0), (3, 0), (2, 20)])
def test_nested_batch_equals_per_cluster(b200_harness, k, gibbs):
    arena = hap_arena(60 + k, n=12)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=k, num_gibbs_samples=gibbs)
    _, one_by_one = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=5, num_threads=1)
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=5)
    for a, b in zip(batched, one_by_one):
        # the sub-EMs run the same kernels on the same rows whether batched or not: identical bits expected
        same(a, b, exact=True)
        assert len(a["gibbs"]) == len(b["gibbs"])
        for ga, gb in zip(a["gibbs"], b["gibbs"]):
            assert np.array_equal(ga["path_ids"], gb["path_ids"])
            assert np.array_equal(ga["noise_samples"], gb["noise_samples"]) and np.array_equal(ga["abundance_samples"], gb["abundance_samples"])


def test_nested_batch_matches_reference_class(b200_harness, reference):
    arena = hap_arena(71, n=10)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena)
    for a, b in zip(batched, ref):
        assert a["sets"] == b["sets"]
        assert np.max(np.abs(a["posteriors"] - b["posteriors"]), initial=0) < 1e-9
        scale = max(1.0, b["total_count"])
        assert np.max(np.abs(a["abundances"] - b["abundances"]), initial=0) <= 1e-6 * scale
        assert abs(a["noise_count"] - b["noise_count"]) <= 1e-6 * scale


@pytest.mark.parametrize("gibbs", [0, 10])
def test_transcripts_batch_equals_per_cluster(b200_harness, gibbs):
    arena = synth.transcripts_arena(seed=31, n_paths=800, n_reads=4000, n_max=90)
    cfg = loader.make_config(loader.MODEL_TRANSCRIPTS, num_gibbs_samples=gibbs)
    _, one_by_one = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=3, num_threads=1)
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=3)
    for a, b in zip(batched, one_by_one):
        same(a, b, exact=True)
        assert len(a["gibbs"]) == len(b["gibbs"])
        for ga, gb in zip(a["gibbs"], b["gibbs"]):
            assert np.array_equal(ga["noise_samples"], gb["noise_samples"]) and np.array_equal(ga["abundance_samples"], gb["abundance_samples"])


def test_other_models_fall_back_to_the_loop(b200_harness):
    arena = hap_arena(80, n=4)
    for cfg in (loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2, infer_collapsed=False, min_hap_prob=0.01),
                loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2), loader.make_config(loader.MODEL_STRAINS)):
        _, one_by_one = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=9, num_threads=1)
        _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=9)
        for a, b in zip(batched, one_by_one):
            same(a, b, exact=True)


@pytest.mark.parametrize("model, k", [(loader.MODEL_PATH_POSTERIOR, 1), (loader.MODEL_HAPLOTYPES, 2), (loader.MODEL_HAPLOTYPES, 3), (loader.MODEL_HAPLOTYPES, 1)])
def test_posterior_batch_equals_per_cluster_and_reference(b200_harness, reference, model, k):
    """Haplotypes model: raw matrices of all clusters on the device once, one log-likelihood call per round."""
    arena = hap_arena(90 + k, n=9)
    cfg = loader.make_config(model, group_size=k)
    _, one_by_one = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=2, num_threads=1)
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=2)
    for a, b, r in zip(batched, one_by_one, ref):
        same(a, b, exact=True)
        assert a["sets"] == r["sets"]
        assert np.max(np.abs(a["posteriors"] - r["posteriors"]), initial=0) < 1e-9


@pytest.mark.parametrize("gibbs", [0, 8])
def test_strains_batch_equals_per_cluster(b200_harness, gibbs):
    arena = concat([synth.transcripts_arena(seed=33, n_paths=500, n_reads=3000, n_max=70), hap_arena(95, n=5)])
    cfg = loader.make_config(loader.MODEL_STRAINS, num_gibbs_samples=gibbs)
    _, one_by_one = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=4, num_threads=1)
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=4)
    for a, b in zip(batched, one_by_one):
        same(a, b, exact=True)
        assert len(a["gibbs"]) == len(b["gibbs"])
        for ga, gb in zip(a["gibbs"], b["gibbs"]):
            assert np.array_equal(ga["path_ids"], gb["path_ids"])
            assert np.array_equal(ga["noise_samples"], gb["noise_samples"]) and np.array_equal(ga["abundance_samples"], gb["abundance_samples"])


def test_posterior_batch_on_single_path_and_mixed_clusters(b200_harness, reference):
    """Diploid posteriors on clusters of 1 to ~60 paths (a single path has a single pair): the device pair walk and the
    per-cluster path agree exactly, and with the reference class."""
    arena = synth.transcripts_arena(seed=37, n_paths=600, n_reads=3500, n_max=60)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2)
    _, one_by_one = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=6, num_threads=1)
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=6)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=6)
    assert min(len(r["sets"]) for r in ref if r["total_count"] > 0 or r["sets"]) >= 1
    for a, b, r in zip(batched, one_by_one, ref):
        same(a, b, exact=True)
        assert a["sets"] == r["sets"]
        assert np.max(np.abs(a["posteriors"] - r["posteriors"]), initial=0) < 1e-9


def test_pinned_scratch(ctx):
    """rpvg_b200_pinned_scratch: a slot keeps its buffer while requests fit, grows otherwise, slots are independent,
    the memory is writable host memory, and a slot outside 0..7 is refused."""
    import ctypes as C

    a = ctx.pinned_scratch(0, 1000)
    assert ctx.pinned_scratch(0, 500) == a
    assert ctx.pinned_scratch(0, 1000) == a
    b = ctx.pinned_scratch(1, 1 << 20)
    assert b != a

    view = (C.c_ubyte * (1 << 20)).from_address(b)
    view[0] = 7
    view[(1 << 20) - 1] = 9
    assert view[0] == 7 and view[(1 << 20) - 1] == 9

    big = ctx.pinned_scratch(0, 8 << 20)     # beyond the first allocation and its headroom
    view = (C.c_ubyte * (8 << 20)).from_address(big)
    view[(8 << 20) - 1] = 1
    assert ctx.pinned_scratch(0, 1000) == big
    assert ctx.pinned_scratch(1, 1000) == b

    with pytest.raises(capi.RpvgB200Error):
        ctx.pinned_scratch(8, 16)
