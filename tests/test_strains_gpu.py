"""Strains model on the GPU: greedy weighted path cover (rpvg_b200_path_cover) and MinimumPathAbundanceEstimator, against
the reference's own code (oracle/_ref) and its frozen output (tests/golden). The cover is integer work: it must be
identical; abundances within 1e-6 relative."""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth
from rpvg_b200.arena import ClusterArena, concat
from tests.golden_util import cluster_results, load

pytestmark = pytest.mark.gpu

REL_TOL = 1e-6


def reference_cover(ref, arena, c):
    """The cover the reference builds inside MinimumPathAbundanceEstimator::estimate (path_abundance_estimator.cpp:229-252)."""
    mat, noise, counts = arena.dense(c)
    rows, n = mat.shape
    counts = counts.copy()
    counts[np.isclose(noise, 1.0, rtol=2.3e-14, atol=0)] = 0
    cover = np.asfortranarray((mat > 0).astype(np.uint8))
    with np.errstate(divide="ignore"):
        logs = np.where(mat > 0, np.log(np.where(mat > 0, mat, 1.0)), 0.0)
    weights = np.zeros(n)
    for i in range(rows):            # row order, like the reference's accumulation
        weights += logs[i] * counts[i]
    weights *= -1
    out = np.zeros(n, dtype=np.uint32)
    k = ref.ref_path_cover(rows, n, cover.ctypes.data, counts.ctypes.data, weights.ctypes.data, out.ctypes.data)
    flags = np.zeros(n, dtype=bool)
    flags[out[:k]] = True
    return flags


def redundant_arena(seed):
    """Config-5 style: transcript clusters with ~30% of the paths duplicated (copied columns with perturbed values)."""
    base = synth.transcripts_arena(seed=seed, n_paths=500, n_reads=3000, n_max=80)
    rng = np.random.default_rng(seed)
    read_offset, path_offset, row_ptr, col, prob = [0], [0], [0], [], []
    for c in range(base.n_clusters):
        M, N, _ = base.cluster_shape(c)
        dup = {j: N + k for k, j in enumerate(np.nonzero(rng.random(N) < 0.3)[0])}
        r0 = int(base.read_offset[c])
        for i in range(M):
            e0, e1 = int(base.row_ptr[r0 + i]), int(base.row_ptr[r0 + i + 1])
            entries = [(int(j), float(p)) for j, p in zip(base.col[e0:e1], base.prob[e0:e1])]
            entries += [(dup[j], p * float(rng.uniform(0.5, 1.0))) for j, p in entries if j in dup and rng.random() < 0.8]
            entries.sort()
            col += [j for j, _ in entries]
            prob += [p for _, p in entries]
            row_ptr.append(len(col))
        read_offset.append(read_offset[-1] + M)
        path_offset.append(path_offset[-1] + N + len(dup))
    return ClusterArena(np.array(read_offset), np.array(path_offset), np.array(row_ptr), np.array(col), np.array(prob), base.noise, base.count)


@pytest.mark.parametrize("seed", [1, 2])
def test_path_cover_is_identical_to_reference(ctx, reference, seed):
    arena = redundant_arena(seed)
    flags = ctx.path_cover(arena)
    sizes = []
    for c in range(arena.n_clusters):
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        expect = reference_cover(reference, arena, c)
        assert np.array_equal(flags[p0:p1], expect), f"cluster {c}"
        sizes.append((int(expect.sum()), p1 - p0))
    assert any(k < n for k, n in sizes if n > 2), "the workload must exercise a non-trivial cover"


def test_path_cover_known_answer(ctx):
    """rpvg:src/tests/path_abundance_estimator_test.cpp:8-28 as an arena: weights from probabilities exp(-w_j / rows_j)."""
    rows = [[0], [0, 1], [1, 2], [1]]
    for probs, expect in [({0: np.exp(-0.5), 1: np.exp(-1 / 3), 2: np.exp(-1.0)}, [True, True, False]),
                          ({0: np.exp(-0.5), 1: np.exp(-1 / 3), 2: np.exp(-0.01)}, [True, True, True])]:
        col = [j for r in rows for j in r]
        prob = [probs[j] for r in rows for j in r]
        arena = ClusterArena(np.array([0, 4]), np.array([0, 3]), np.cumsum([0] + [len(r) for r in rows]), np.array(col), np.array(prob),
                             np.full(4, 0.01), np.ones(4, dtype=np.uint32))
        assert list(ctx.path_cover(arena)) == expect


def test_strains_class_matches_reference_class(b200_harness, reference):
    arena = redundant_arena(3).select(range(40))
    cfg = loader.make_config(loader.MODEL_STRAINS)
    _, mine = loader.harness_estimate_batch(b200_harness, cfg, arena, num_threads=2)
    _, ref = loader.harness_estimate_batch(reference, cfg, arena)
    for a, b in zip(mine, ref):
        assert a["sets"] == b["sets"]
        assert np.array_equal(a["abundances"] == 0, b["abundances"] == 0)
        nz = b["abundances"] != 0
        if nz.any():
            assert np.max(np.abs(a["abundances"][nz] - b["abundances"][nz]) / b["abundances"][nz]) < REL_TOL
        assert abs(a["noise_count"] - b["noise_count"]) <= REL_TOL * max(1.0, b["total_count"])
        assert a["total_count"] == b["total_count"]


def test_strains_class_matches_golden(b200_harness):
    arena, z = load("haplotype_models.npz")
    _, results = loader.harness_estimate_batch(b200_harness, loader.make_config(loader.MODEL_STRAINS), arena, rng_seed=7, num_threads=2)
    for c, res in enumerate(results):
        sets, post, ab, noise, total = cluster_results(z, "strains_", c)
        assert res["sets"] == sets
        assert np.array_equal(res["abundances"] == 0, ab == 0)
        assert np.allclose(res["abundances"], ab, rtol=REL_TOL, atol=0)
        assert abs(res["noise_count"] - noise) <= REL_TOL * max(1.0, total) and res["total_count"] == total


def test_public_weighted_minimum_path_cover_known_answers(ctx):
    """The reference's own test of the hot path (rpvg:src/tests/path_abundance_estimator_test.cpp:8-28) through the C entry
    point behind the public method MinimumPathAbundanceEstimator::weightedMinimumPathCover."""
    cover = np.array([[1, 0, 1], [0, 1, 0], [1, 0, 0], [0, 1, 1]], dtype=bool)
    counts = [1, 3, 1, 5]
    assert ctx.weighted_path_cover(cover, counts, [1, 1, 1]) == [0, 1]
    assert ctx.weighted_path_cover(cover, counts, [1, 1, 0.01]) == [0, 1, 2]
    assert ctx.weighted_path_cover(np.ones((3, 1), dtype=bool), [1, 1, 1], [2.0]) == [0]


def test_reference_catch2_test_compiled_against_the_drop_in_classes():
    """oracle/_ref/dropin_catch_tests: rpvg's own path_abundance_estimator_test.cpp compiled against this repository's
    classes inside rpvg's type system (-DRPVG_B200_WITH_RPVG_HEADERS, built by oracle/Makefile where the reference tree
    exists; the binary travels to the GPU box)."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "dropin_catch_tests")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/dropin_catch_tests is not built")
    out = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "All tests passed" in out.stdout
