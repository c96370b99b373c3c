"""Parity of the CUDA EM path (through the C ABI) with the CPU oracle, transcripts model.

Tolerances are north_star's: per-path read counts within 1e-6 relative, identical set of paths above the min-abundance
threshold; here additionally identical EM iteration counts (same stopping iteration).
"""

import numpy as np
import pytest

from rpvg_b200 import capi, synth
from rpvg_b200.arena import ClusterArena, concat
from oracle import loader

pytestmark = pytest.mark.gpu

REL_TOL = 1e-6


def check_against_port(ctx, arena, params=None, **port_kwargs):
    abundances, noise_count, total_count, iterations = ctx.em_abundances(arena, params)
    o_ab, o_noise, o_total, o_its, _ = loader.port_em_abundances(arena, **port_kwargs)

    assert np.array_equal(total_count, o_total)
    assert np.array_equal(abundances == 0, o_ab == 0), "set of paths above min_em_abundance differs"
    nz = o_ab != 0
    if nz.any():
        rel = np.max(np.abs(abundances[nz] - o_ab[nz]) / o_ab[nz])
        assert rel < REL_TOL, rel
    assert np.allclose(noise_count, o_noise, rtol=REL_TOL, atol=1e-9)
    flips = int(np.sum(iterations != o_its))
    assert flips == 0, f"{flips} clusters stopped on a different EM iteration than the oracle"
    # writer invariant (rpvg:src/threaded_output_writer.cpp:327-328): abundance_sum + noise_count ~ total_count
    for c in range(arena.n_clusters):
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        if total_count[c] > 0:
            assert abs(abundances[p0:p1].sum() + noise_count[c] - total_count[c]) <= 1e-9 * total_count[c]
    return abundances, noise_count, total_count, iterations


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_em_matches_oracle_small_clusters(ctx, seed):
    arena = synth.transcripts_arena(seed=seed, n_paths=3000, n_reads=15000, n_max=300)
    check_against_port(ctx, arena)


def test_em_matches_oracle_mixed_sizes(ctx):
    arena = synth.transcripts_arena(seed=5, n_paths=20000, n_reads=100000, n_max=2000)
    check_against_port(ctx, arena)


def test_em_matches_reference_estimator(ctx, reference):
    """Same clusters through the reference's own PathAbundanceEstimator::estimate (oracle/_ref)."""
    arena = synth.transcripts_arena(seed=9, n_paths=1500, n_reads=8000, n_max=150)
    abundances, noise_count, total_count, _ = ctx.em_abundances(arena)
    _, results = loader.harness_estimate_batch(reference, loader.make_config(loader.MODEL_TRANSCRIPTS), arena)
    for c, res in enumerate(results):
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        mine, ref = abundances[p0:p1], res["abundances"]
        assert np.array_equal(mine == 0, ref == 0)
        nz = ref != 0
        if nz.any():
            assert np.max(np.abs(mine[nz] - ref[nz]) / ref[nz]) < REL_TOL
        assert abs(noise_count[c] - res["noise_count"]) <= REL_TOL * max(1.0, res["total_count"])
        assert total_count[c] == res["total_count"]
        assert res["sets"] == [(j,) for j in range(p1 - p0)]


def test_em_parameters_are_honoured(ctx):
    arena = synth.transcripts_arena(seed=4, n_paths=1000, n_reads=5000, n_max=100)
    params = capi.default_em_params(max_em_its=7, max_rel_em_conv=0.01)
    _, _, _, its = check_against_port(ctx, arena, params, max_em_its=7, max_rel_em_conv=0.01)
    assert its.max() <= 7


def test_edge_cases(ctx):
    """Empty cluster (no read classes), noise-only read classes, single path, single read class."""
    read_offset = np.array([0, 0, 2, 3, 5], dtype=np.uint64)
    path_offset = np.array([0, 3, 5, 6, 8], dtype=np.uint64)
    # cluster 0: empty; cluster 1: one noise-only row + one normal row; cluster 2: single path; cluster 3: two rows
    row_ptr = np.array([0, 0, 2, 3, 4, 6], dtype=np.uint64)
    col = np.array([0, 1, 0, 0, 0, 1], dtype=np.uint32)
    prob = np.array([0.3, 0.6, 0.9, 0.5, 0.2, 0.7], dtype=np.float64)
    noise = np.array([1.0, 0.1, 0.1, 0.5, 0.1], dtype=np.float64)
    count = np.array([3, 5, 7, 1, 4000000000], dtype=np.uint32)
    arena = ClusterArena(read_offset, path_offset, row_ptr, col, prob, noise, count)
    abundances, noise_count, total_count, iterations = check_against_port(ctx, arena)
    assert iterations[0] == 0 and total_count[0] == 0 and np.all(abundances[:3] == 0)
    assert total_count[3] == 4000000001.0


def test_malformed_arena_is_rejected(ctx):
    arena = synth.transcripts_arena(seed=4, n_paths=200, n_reads=1000, n_max=50)
    bad = ClusterArena(arena.read_offset, arena.path_offset, arena.row_ptr, arena.col.copy(), arena.prob, arena.noise, arena.count)
    bad.col[0] = 10 ** 6
    with pytest.raises(capi.RpvgB200Error, match="out of range"):
        ctx.em_abundances(bad)
    # the context stays usable
    check_against_port(ctx, arena)


def test_full_size_properties(ctx):
    """Config 2 at full size (1M read classes, 200k paths): properties that need no oracle run —
    mass conservation per cluster, determinism (two runs bit-identical), and permutation of clusters."""
    arena = synth.transcripts_arena()
    a1 = ctx.em_abundances(arena)
    a2 = ctx.em_abundances(arena)
    for x, y in zip(a1, a2):
        assert np.array_equal(x, y), "EM is not run-to-run deterministic"
    abundances, noise_count, total_count, iterations = a1
    sums = np.add.reduceat(abundances, arena.path_offset[:-1].astype(np.int64))
    assert np.allclose(sums + noise_count, total_count, rtol=1e-9)
    counts = np.add.reduceat(arena.count.astype(np.float64), arena.read_offset[:-1].astype(np.int64))
    assert np.array_equal(counts, total_count)
    assert iterations.min() >= 10 and iterations.max() <= 10000
    # cluster order must not matter: reverse the batch
    rev = arena.select(list(range(arena.n_clusters - 1, -1, -1))[:500])
    r_ab, r_noise, _, r_its = ctx.em_abundances(rev)
    for k in range(500):
        c = arena.n_clusters - 1 - k
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        q0, q1 = int(rev.path_offset[k]), int(rev.path_offset[k + 1])
        assert np.array_equal(r_ab[q0:q1], abundances[p0:p1])
        assert r_its[k] == iterations[c]


def _random_cluster(rng, m, n, per_row, dense=False):
    """One cluster as arena pieces: m read classes over n paths, per_row paths each (all n when dense)."""
    if dense:
        cols = np.tile(np.arange(n, dtype=np.uint32), m)
        lens = np.full(m, n, dtype=np.int64)
    else:
        lens = rng.integers(1, 2 * per_row, size=m)
        lens = np.minimum(lens, n)
        cols = np.concatenate([np.sort(rng.choice(n, size=k, replace=False)).astype(np.uint32) for k in lens])
    row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    prob = rng.uniform(1e-6, 1.0, size=cols.size)
    noise = rng.uniform(1e-4, 0.05, size=m)
    count = rng.integers(1, 50, size=m).astype(np.uint32)
    return ClusterArena(np.array([0, m], dtype=np.uint64), np.array([0, n], dtype=np.uint64), row_ptr, cols, prob, noise, count)


def test_em_groups_without_band_structure(ctx):
    """Clusters that are split over several CTAs and in which every slice touches (nearly) every path: the exchange of
    partial column sums between the CTAs is all-to-all then (the band structure config 2 has is what makes it cheap,
    not what makes it correct). Dense small-N, dense mid-N and scattered large-N clusters, next to a tiny one."""
    rng = np.random.default_rng(77)
    parts = [
        _random_cluster(rng, 1500, 24, 0, dense=True),       # weight 37.5k -> group of 4
        _random_cluster(rng, 4000, 40, 0, dense=True),       # weight 164k -> group of 16, every column in every slice
        _random_cluster(rng, 20000, 3000, 5),                # scattered: ~100k entries over 3000 paths, group of 8-16
        _random_cluster(rng, 30, 4, 2),
        _random_cluster(rng, 9000, 700, 3),                  # group of 2-4
    ]
    arena = concat(parts)
    _, _, _, its = check_against_port(ctx, arena)
    assert its.min() >= 10


def test_em_wide_clusters(ctx):
    """Clusters with more than 65535 paths (32-bit indices in the kernels): one split over a group of CTAs, one small
    enough for a single CTA but with an abundance vector that does not fit shared memory (global-memory fallback)."""
    rng = np.random.default_rng(91)
    parts = [
        _random_cluster(rng, 30000, 70000, 3),    # weight ~120k -> group, each slice touches a few thousand paths
        _random_cluster(rng, 800, 90000, 4),      # one CTA, N + 1 doubles do not fit shared memory
        _random_cluster(rng, 50, 6, 2),
    ]
    arena = concat(parts)
    check_against_port(ctx, arena)


def test_em_dense_kernels(ctx, monkeypatch):
    """Small clusters run on the dense kernels (em_dense.cuh; max(read classes, paths + 1) <= 32 by default): sizes on
    both sides of every kernel side, full matrices, a path nobody maps to, a cluster without read classes — against the
    oracle, and against the sliced-ELL kernels on the same arena (RPVG_B200_EM_DENSE_MAX=0), also with every side the
    library has (128)."""
    rng = np.random.default_rng(123)
    parts = []
    for m, n in [(1, 1), (2, 1), (15, 15), (16, 15), (16, 16), (17, 3), (3, 17), (31, 31), (32, 31), (32, 32), (33, 31), (40, 12),
                 (63, 63), (64, 63), (65, 20), (100, 127), (128, 127), (128, 128), (129, 5)]:
        parts.append(_random_cluster(rng, m, n, max(1, n // 3)))
    parts.append(_random_cluster(rng, 16, 15, 0, dense=True))
    parts.append(_random_cluster(rng, 32, 31, 0, dense=True))
    parts.append(_random_cluster(rng, 128, 127, 0, dense=True))

    # a path without entries (column 2 removed from every read class) and a cluster without read classes
    holed = _random_cluster(rng, 12, 6, 2)
    keep = holed.col != 2
    lens = np.add.reduceat(keep.astype(np.int64), holed.row_ptr[:-1].astype(np.int64))
    row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    parts.append(ClusterArena(holed.read_offset, holed.path_offset, row_ptr, holed.col[keep], holed.prob[keep], holed.noise, holed.count))
    parts.append(ClusterArena(np.array([0, 0], dtype=np.uint64), np.array([0, 3], dtype=np.uint64), np.array([0], dtype=np.uint64), np.zeros(0, dtype=np.uint32),
                              np.zeros(0), np.zeros(0), np.zeros(0, dtype=np.uint32)))

    arena = concat(parts)

    results = {}
    for dense_max in ("0", "32", "128"):
        monkeypatch.setenv("RPVG_B200_EM_DENSE_MAX", dense_max)
        results[dense_max] = check_against_port(ctx, arena)

    for dense_max in ("32", "128"):
        for mine, sell in zip(results[dense_max], results["0"]):
            assert np.allclose(mine, sell, rtol=REL_TOL, atol=1e-9)
        assert np.array_equal(results[dense_max][3], results["0"][3])


def giant_arena(n_paths_giant, n_reads, seed=5):
    """One giant cluster and a handful of ordinary ones around it."""
    sizes = [7, 300, n_paths_giant, 40, 1200]
    return synth.transcripts_arena(seed=seed, n_reads=n_reads, sizes=sizes, reads_cap=1e9)


def test_em_giant_cluster_streams_over_the_whole_gpu(ctx):
    """A cluster far beyond 16 shared-memory slices goes to emStreamKernel (one cooperative launch of the whole GPU, the
    matrix streamed every iteration, grid-wide barriers). To convergence against the oracle: abundances, zero set and the
    stopping iteration."""
    arena = giant_arena(8000, 100_000)
    shapes = [arena.cluster_shape(c) for c in range(arena.n_clusters)]
    weight = max(m + z for m, _, z in shapes)
    assert weight > 500_000

    ab, noise, total, its = ctx.em_abundances(arena)
    o_ab, o_noise, o_total, o_its, _ = loader.port_em_abundances(arena, num_threads=0)

    assert np.array_equal(total, o_total)
    assert np.array_equal(ab == 0, o_ab == 0)
    nz = o_ab != 0
    assert np.max(np.abs(ab[nz] - o_ab[nz]) / o_ab[nz]) < 1e-6
    assert np.allclose(noise, o_noise, rtol=1e-6, atol=1e-9)
    assert np.array_equal(its, o_its)

    # bit-reproducible
    ab2, noise2, _, its2 = ctx.em_abundances(arena)
    assert np.array_equal(ab, ab2) and np.array_equal(noise, noise2) and np.array_equal(its, its2)


def test_em_giant_cluster_two_million_entries(ctx):
    """More than 2 M entries in one cluster (48 MB in the two orientations), 250 iterations on both sides."""
    arena = giant_arena(40000, 520_000, seed=6)
    assert max(z for _, _, z in (arena.cluster_shape(c) for c in range(arena.n_clusters))) > 2_000_000

    params = capi.default_em_params(max_em_its=250)
    ab, noise, total, its = ctx.em_abundances(arena, params)
    o_ab, o_noise, o_total, o_its, _ = loader.port_em_abundances(arena, max_em_its=250, num_threads=0)

    assert np.array_equal(its, o_its) and int(its.max()) == 250
    assert np.array_equal(ab == 0, o_ab == 0)
    nz = o_ab != 0
    assert np.max(np.abs(ab[nz] - o_ab[nz]) / o_ab[nz]) < 1e-6
    assert np.allclose(noise, o_noise, rtol=1e-6, atol=1e-9)
