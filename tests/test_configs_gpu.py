"""Parity at BASELINE.json's own configurations (generators in rpvg_b200/synth.py):

  C2  the headline arena at FULL size against the CPU oracle: every cluster, abundances 1e-6, identical zero sets,
      identical EM iteration counts;
  C1  a slice of the example-structured arena through NestedPathAbundanceEstimator against the reference's class;
  C3  one 1000GP-style cluster with >= 300 haplotype groups and >= 20,000 read classes through the haplotype-transcripts
      class, posteriors 1e-9, tuples in the reference's order;
  C4  the haplotypes model: diploid bounded posteriors on >= 300 paths x >= 20,000 read classes, and --use-hap-gibbs at
      ploidy 4 on >= 200 paths (the chain replays the reference's with the same mt19937);
  C5  the strains model: the greedy cover of a cluster with >= 2,000 paths identical to the reference's rule, and the
      whole class against the reference's on a cluster with > 800 paths.
The reference side is rpvg's own classes (oracle/_ref), one thread per cluster like rpvg.
"""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth

pytestmark = pytest.mark.gpu


def compare_estimates(mine, ref, posterior_tol=1e-9, rel=1e-6):
    assert mine["sets"] == ref["sets"], "path tuples / their order differ"
    assert np.max(np.abs(mine["posteriors"] - ref["posteriors"]), initial=0) < posterior_tol
    assert len(mine["abundances"]) == len(ref["abundances"])
    scale = max(1.0, ref["total_count"])
    assert np.max(np.abs(mine["abundances"] - ref["abundances"]), initial=0) <= rel * scale
    big = ref["abundances"] > 1e-3 * scale
    if big.any():
        assert np.max(np.abs(mine["abundances"][big] - ref["abundances"][big]) / ref["abundances"][big]) < rel
    assert abs(mine["noise_count"] - ref["noise_count"]) <= rel * scale
    assert mine["total_count"] == ref["total_count"]


def single_transcript_cluster(n_paths, n_reads, seed, ploidy=2):
    rng = np.random.default_rng(seed)
    st = synth.synthetic_structure(rng, 1, n_haplotypes=5008, paths_per_transcript=np.array([n_paths]), zipf=0.6)
    return synth.haplotype_arena(st, n_reads, seed=seed + 1, ploidy=ploidy, single_transcript_clusters=True, cross_transcript=0.0)


# ---- C2 -----------------------------------------------------------------------------------------------------------------

def test_config2_full_size_against_the_oracle(ctx):
    arena = synth.transcripts_arena()    # 1M read classes, 200k paths, cluster sizes 1-5000: the bench workload itself
    assert arena.n_reads > 1_000_000 - 1000 and arena.n_paths == 200_000

    abundances, noise_count, total_count, iterations = ctx.em_abundances(arena)
    o_ab, o_noise, o_total, o_its, _ = loader.port_em_abundances(arena, num_threads=0)

    assert np.array_equal(total_count, o_total)
    assert np.array_equal(abundances == 0, o_ab == 0), "set of paths above min_em_abundance differs"
    nz = o_ab != 0
    assert np.max(np.abs(abundances[nz] - o_ab[nz]) / o_ab[nz]) < 1e-6
    assert np.allclose(noise_count, o_noise, rtol=1e-6, atol=1e-9)
    assert np.array_equal(iterations, o_its), f"EM iteration counts differ in {int((iterations != o_its).sum())} of {arena.n_clusters} clusters"

    # the clusters that run on 16-CTA groups are in there
    sizes = np.diff(arena.path_offset.astype(np.int64))
    assert sizes.max() > 4000 and int(o_its.max()) == 10000


# ---- C1 -----------------------------------------------------------------------------------------------------------------

def test_config1_example_slice_nested_model(b200_harness, reference):
    st = synth.example_structure()
    sizes = np.diff(st.transcript_ptr)
    # transcripts of moderate size, so that the reference's dense pair loop stays within seconds per cluster
    pick = np.nonzero((sizes >= 2) & (sizes <= 40))[0][:120]
    arena = synth.haplotype_arena(st, 12000, seed=synth.BASE_SEED + 1, transcripts=pick)
    assert arena.n_clusters > 40 and arena.n_reads > 3000

    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=5, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=5)
    n_sets = 0
    for a, b in zip(mine, ref):
        compare_estimates(a, b)
        n_sets += len(b["sets"])
    assert n_sets > arena.n_clusters

    # the batched driver gives the same
    _, batched = loader.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=5)
    for a, b in zip(batched, ref):
        compare_estimates(a, b)


# ---- C3 / C4 ------------------------------------------------------------------------------------------------------------

@pytest.fixture(scope="module")
def big_cluster():
    arena = single_transcript_cluster(340, 300_000, seed=5)
    assert arena.n_paths >= 300 and arena.n_reads >= 20_000
    return arena


def test_config3_large_cluster_nested_model(b200_harness, reference, big_cluster):
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, big_cluster, rng_seed=2, num_threads=1)
    _, ref = loader.harness_estimate_batch(reference, cfg, big_cluster, rng_seed=2, num_threads=1)
    compare_estimates(mine[0], ref[0])


def test_config3_many_diplotypes_nested_model(b200_harness, reference):
    """Few reads on many haplotype groups: hundreds of diplotypes stay above min_hap_prob, each with its own sub-EM."""
    arena = single_transcript_cluster(300, 15, seed=8)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=2, num_threads=1)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=2, num_threads=1)
    assert len(ref[0]["sets"]) > 100
    compare_estimates(mine[0], ref[0])


def test_config4_diploid_bounded_posteriors(b200_harness, reference, big_cluster):
    cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, big_cluster, num_threads=1)
    _, ref = loader.harness_estimate_batch(reference, cfg, big_cluster, num_threads=1)
    compare_estimates(mine[0], ref[0])


def test_config4_diploid_bounded_posteriors_flat_likelihood(b200_harness, reference):
    """Few reads: thousands of pairs survive the bound, the list and its order must still be the reference's."""
    arena = single_transcript_cluster(320, 8, seed=12)
    cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, num_threads=1)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, num_threads=1)
    assert len(ref[0]["sets"]) > 10000
    compare_estimates(mine[0], ref[0])


def test_config4_ploidy4_haplotype_gibbs(b200_harness, reference):
    arena = single_transcript_cluster(215, 4000, seed=6, ploidy=4)
    assert arena.n_paths >= 200
    cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=4, use_group_post_gibbs=True)
    for seed in (3, 4):
        _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, rng_seed=seed, num_threads=1)
        _, ref = loader.harness_estimate_batch(reference, cfg, arena, rng_seed=seed, num_threads=1)
        compare_estimates(mine[0], ref[0])   # same generator, same chain: same visit frequencies


# ---- C5 -----------------------------------------------------------------------------------------------------------------

def reference_rule_cover(arena, c):
    """weightedMinimumPathCover (rpvg:src/path_abundance_estimator.cpp:229-251, 297-340) on the non-zeros. The scores are
    quotients of exact integer sums and the weights, so the order of the additions does not matter."""
    M, N, _ = arena.cluster_shape(c)
    r0 = int(arena.read_offset[c])
    e0, e1 = int(arena.row_ptr[r0]), int(arena.row_ptr[r0 + M])
    rows = np.repeat(np.arange(M), np.diff(arena.row_ptr[r0:r0 + M + 1].astype(np.int64)))
    cols = arena.col[e0:e1].astype(np.int64)
    counts = arena.count[r0:r0 + M].astype(np.float64)
    noise = arena.noise[r0:r0 + M]
    counts[(noise == 1) | (np.abs(noise - 1) < np.minimum(noise, 1) * 100 * np.finfo(np.float64).eps)] = 0
    weights = np.zeros(N)
    np.add.at(weights, cols, np.log(arena.prob[e0:e1]) * counts[rows])
    weights *= -1
    if N == 1:
        return [0]
    uncovered = counts.copy()
    cover = []
    order = np.argsort(cols, kind="stable")
    col_ptr = np.searchsorted(cols[order], np.arange(N + 1))
    with np.errstate(divide="ignore", invalid="ignore"):
        while uncovered.max() > 0:
            score = np.bincount(cols, weights=uncovered[rows], minlength=N) / weights
            best = int(np.argmax(np.where(np.isnan(score), -1, score)))   # first strict maximum
            assert score[best] > 0
            cover.append(best)
            uncovered[rows[order[col_ptr[best]:col_ptr[best + 1]]]] = 0
    return sorted(cover)


def test_config5_strains_cover_and_class(ctx, b200_harness, reference):
    arena = synth.strains_arena(n_paths=40_000, n_reads=200_000, n_max=5000)
    sizes = np.diff(arena.path_offset.astype(np.int64))
    big = int(np.argmax(sizes))
    assert sizes[big] >= 2000

    # the cover of every cluster in one call; the rule's cover for the biggest clusters
    in_cover = ctx.path_cover(arena)
    for c in np.argsort(-sizes)[:4]:
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        want = reference_rule_cover(arena, int(c))
        assert list(np.nonzero(in_cover[p0:p1])[0]) == want
        assert 0 < len(want) < sizes[c]          # the duplicated paths make the cover a proper subset

    # the whole class against the reference's class on a cluster with > 800 paths (the reference is dense: ~20 s there)
    mid = int(np.argmin(np.abs(sizes - 900)))
    assert sizes[mid] > 800
    cfg = loader.make_config(loader.MODEL_STRAINS)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, [mid], num_threads=1)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena, [mid], num_threads=1)
    assert np.array_equal(mine[0]["abundances"] == 0, ref[0]["abundances"] == 0)
    compare_estimates(mine[0], ref[0])
    p0, p1 = int(arena.path_offset[mid]), int(arena.path_offset[mid + 1])
    assert list(np.nonzero(in_cover[p0:p1])[0]) == reference_rule_cover(arena, mid)
    assert set(np.nonzero(ref[0]["abundances"])[0]) <= set(np.nonzero(in_cover[p0:p1])[0])
