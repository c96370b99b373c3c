"""Gibbs read-count sampler on the device (Philox counter-based RNG) against the reference's sampler (mt19937 +
libstdc++ distributions) — agreement in distribution over fixed seeds (north_star), exact invariants per sample, and
bit-reproducibility for equal seeds."""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth
from rpvg_b200.arena import concat

pytestmark = pytest.mark.gpu


def small_arena(seed, n=3):
    rng = np.random.default_rng(seed)
    return concat([synth.haplotype_cluster(rng, n_transcripts=2, paths_per_transcript=(2, 4), n_haplotypes=12, n_reads=int(rng.integers(150, 400)))
                   for _ in range(n)])


def moments_agree(a, b, n_eff):
    """Per column: means within 6 standard errors (samples are autocorrelated: n_eff < n), spreads within a factor."""
    ma, mb = a.mean(axis=0), b.mean(axis=0)
    sa, sb = a.std(axis=0), b.std(axis=0)
    se = np.sqrt((sa ** 2 + sb ** 2) / n_eff) + 1e-9
    assert np.all(np.abs(ma - mb) <= 6 * se + 1e-6 * np.maximum(1, np.abs(mb))), (ma, mb, se)
    big = (sa > 1e-3) & (sb > 1e-3)
    if big.any():
        ratio = sa[big] / sb[big]
        assert np.all((ratio > 0.6) & (ratio < 1.6)), ratio


def test_transcripts_gibbs_matches_reference_in_distribution(b200_harness, reference):
    arena = small_arena(3)
    samples = 1500
    cfg = loader.make_config(loader.MODEL_TRANSCRIPTS, num_gibbs_samples=samples, gibbs_thin_its=3)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=1, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=1)
    for a, b in zip(mine, ref):
        assert len(a["gibbs"]) == 1 and len(b["gibbs"]) == 1
        ga, gb = a["gibbs"][0], b["gibbs"][0]
        assert np.array_equal(ga["path_ids"], gb["path_ids"])
        assert ga["abundance_samples"].shape == gb["abundance_samples"].shape == (samples, len(ga["path_ids"]))
        # every sample conserves the read count (rpvg:src/threaded_output_writer.cpp invariant)
        assert np.allclose(ga["abundance_samples"].sum(axis=1) + ga["noise_samples"], a["total_count"], rtol=1e-9)
        moments_agree(ga["abundance_samples"], gb["abundance_samples"], n_eff=samples / 4)
        moments_agree(ga["noise_samples"][:, None], gb["noise_samples"][:, None], n_eff=samples / 4)
        # the sampler scatters around the EM estimate it starts from
        assert np.all(np.abs(ga["abundance_samples"].mean(axis=0) - a["abundances"]) <= 0.25 * a["total_count"])


def test_gibbs_is_reproducible_and_seed_dependent(b200_harness):
    arena = small_arena(5, n=2)
    cfg = loader.make_config(loader.MODEL_TRANSCRIPTS, num_gibbs_samples=50, gibbs_thin_its=2)
    _, r1 = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=9, num_threads=1)
    _, r2 = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=9, num_threads=2)
    _, r3 = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=10, num_threads=1)
    for a, b, c in zip(r1, r2, r3):
        assert np.array_equal(a["gibbs"][0]["abundance_samples"], b["gibbs"][0]["abundance_samples"])
        assert not np.array_equal(a["gibbs"][0]["abundance_samples"], c["gibbs"][0]["abundance_samples"])


def test_strains_and_nested_gibbs_shapes_and_moments(b200_harness, reference):
    arena = small_arena(8, n=2)
    samples = 600
    for model in (loader.MODEL_STRAINS, loader.MODEL_HAPLOTYPE_TRANSCRIPTS):
        cfg = loader.make_config(model, group_size=2, num_gibbs_samples=samples, gibbs_thin_its=2)
        _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=4, num_threads=2)
        _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=4)
        for a, b in zip(mine, ref):
            assert sum(len(g["noise_samples"]) for g in a["gibbs"]) == samples == sum(len(g["noise_samples"]) for g in b["gibbs"])
            for g in a["gibbs"]:
                assert np.allclose(g["abundance_samples"].sum(axis=1) + g["noise_samples"], a["total_count"], rtol=1e-9)
            if model == loader.MODEL_STRAINS:
                assert np.array_equal(a["gibbs"][0]["path_ids"], b["gibbs"][0]["path_ids"])
                moments_agree(a["gibbs"][0]["abundance_samples"], b["gibbs"][0]["abundance_samples"], n_eff=samples / 4)
            else:
                # The budget is split over the haplotype combinations by sequential host binomials (rpvg:src/path_abundance_estimator.cpp:675-697).
                # The first combination's draw comes from the same generator state in both (exactly equal); after it the
                # reference's sampler has consumed the generator and the device's has not, so the later counts agree as
                # binomial draws with the same parameters do: within 6 standard deviations, and every combination that is
                # expected to get a sizeable share appears on both sides.
                counts_a = {tuple(g["path_ids"]): len(g["noise_samples"]) for g in a["gibbs"]}
                counts_b = {tuple(g["path_ids"]): len(g["noise_samples"]) for g in b["gibbs"]}
                first_a, first_b = tuple(a["gibbs"][0]["path_ids"]), tuple(b["gibbs"][0]["path_ids"])
                assert first_a == first_b and counts_a[first_a] == counts_b[first_b]
                for key in set(counts_a) | set(counts_b):
                    na, nb = counts_a.get(key, 0), counts_b.get(key, 0)
                    share = (na + nb) / (2.0 * samples)
                    assert abs(na - nb) <= 6 * np.sqrt(2 * samples * share * (1 - share)) + 1, (key, na, nb)
                    if share * samples >= 30:
                        assert na > 0 and nb > 0, (key, na, nb)
                for key in set(counts_a) & set(counts_b):
                    if min(counts_a[key], counts_b[key]) >= 200:
                        ga = next(g for g in a["gibbs"] if tuple(g["path_ids"]) == key)
                        gb = next(g for g in b["gibbs"] if tuple(g["path_ids"]) == key)
                        moments_agree(ga["abundance_samples"], gb["abundance_samples"], n_eff=min(counts_a[key], counts_b[key]) / 4)


def test_read_count_sampler_marginals_on_a_wide_cluster(b200_harness, reference):
    """A cluster with more than 50 paths, the default thinning (25 iterations between samples: draws are close to
    independent): per path a two-sample Kolmogorov-Smirnov test between the device's samples and the reference's, and the
    same for the noise count. Bonferroni over the paths at a family-wise 1e-3."""
    from scipy import stats

    arena = synth.transcripts_arena(seed=29, n_paths=3000, n_reads=15000, n_max=120)
    sizes = np.diff(arena.path_offset.astype(np.int64))
    c = int(np.argmax((sizes > 50) & (sizes < 100)))
    assert 50 < sizes[c] < 100
    sub = arena.select([c])

    samples = 800
    cfg = loader.make_config(loader.MODEL_TRANSCRIPTS, num_gibbs_samples=samples, gibbs_thin_its=25)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, sub, rng_seed=2, num_threads=1)
    _, ref = loader.harness_estimate_batch(reference, cfg, sub, rng_seed=2, num_threads=1)
    ga, gb = mine[0]["gibbs"][0], ref[0]["gibbs"][0]
    assert ga["abundance_samples"].shape == gb["abundance_samples"].shape == (samples, sizes[c])

    tested, worst = 0, 1.0
    for j in range(sizes[c]):
        xa, xb = ga["abundance_samples"][:, j], gb["abundance_samples"][:, j]
        if max(xa.std(), xb.std()) < 1e-6:       # a path the data rules out: both stay at (almost) zero
            assert abs(xa.mean() - xb.mean()) < 1e-3 * max(1.0, mine[0]["total_count"])
            continue
        p = stats.ks_2samp(xa, xb).pvalue
        worst = min(worst, p)
        tested += 1
    assert tested > 20
    assert worst > 1e-3 / tested, f"smallest KS p-value {worst} over {tested} paths"
    assert stats.ks_2samp(ga["noise_samples"], gb["noise_samples"]).pvalue > 1e-4
