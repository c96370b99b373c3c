"""rpvg_b200_subset_em (partial matrix -> normalise -> collapse -> EM, batched over path subsets) against the reference's
own protected members called one by one (oracle/_ref: constructPartialProbabilityMatrix, addNoiseAndNormalize...,
readCollapseProbabilityMatrix, EMAbundanceEstimator)."""

import ctypes as C

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth
from rpvg_b200.arena import concat

pytestmark = pytest.mark.gpu

REL_TOL = 1e-6


def reference_subset_em(ref, arena, c, subset):
    data = loader.HarnessClusterData(arena, c)
    M, N, _ = arena.cluster_shape(c)
    ids = np.ascontiguousarray(subset, dtype=np.uint32)
    n = len(ids)
    mat = np.zeros((M, n + 1), order="F"); noise = np.zeros(M); counts = np.zeros(M)
    ref.ref_build_matrix(1e-8, C.byref(data.struct), 1, n, None, ids.ctypes.data, 1, mat.ctypes.data, noise.ctypes.data, counts.ctypes.data)
    out = np.zeros((M, n + 1), order="F"); cnt = np.zeros(M)
    rows = ref.ref_collapse(1e-8, M, n + 1, mat.ctypes.data, counts.ctypes.data, out.ctypes.data, cnt.ctypes.data)
    collapsed = np.asfortranarray(out.ravel(order="F")[:rows * (n + 1)].reshape((rows, n + 1), order="F"))
    ab = np.zeros(n); nc = C.c_double(); tc = C.c_double()
    ref.ref_em(10000, 0.001, rows, n, collapsed.ctypes.data, cnt[:rows].copy().ctypes.data, ab.ctypes.data, C.byref(nc), C.byref(tc))
    return ab, nc.value, tc.value, rows


def test_subset_em_matches_reference_members(ctx, reference):
    rng = np.random.default_rng(17)
    arena = concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 4)), paths_per_transcript=(3, 9), n_haplotypes=30,
                                            n_reads=int(rng.integers(150, 600))) for _ in range(6)])
    subset_cluster, subsets = [], []
    for c in range(arena.n_clusters):
        N = arena.cluster_shape(c)[1]
        for _ in range(5):
            k = int(rng.integers(1, min(N, 6) + 1))
            subsets.append(np.sort(rng.choice(N, size=k, replace=False)))
            subset_cluster.append(c)
        subsets.append(np.arange(N))
        subset_cluster.append(c)

    parts, noise_count, total_count, iterations, n_classes = ctx.subset_em(arena, subset_cluster, subsets)

    for s, (c, subset) in enumerate(zip(subset_cluster, subsets)):
        ab, nc, tc, rows = reference_subset_em(reference, arena, c, subset)
        assert n_classes[s] == rows, "collapse must leave the same number of read classes"
        assert total_count[s] == tc
        assert np.array_equal(parts[s] == 0, ab == 0)
        nz = ab != 0
        if nz.any():
            assert np.max(np.abs(parts[s][nz] - ab[nz]) / ab[nz]) < REL_TOL
        assert abs(noise_count[s] - nc) <= REL_TOL * max(1.0, tc)
        assert abs(parts[s].sum() + noise_count[s] - total_count[s]) <= 1e-9 * total_count[s]


def test_subset_em_on_transcript_clusters(ctx, reference):
    """Power-law clusters (config 2 style): the full-subset EM must equal the plain EM entry point after collapse."""
    arena = synth.transcripts_arena(seed=23, n_paths=400, n_reads=2500, n_max=60)
    rng = np.random.default_rng(5)
    ids = [c for c in range(arena.n_clusters) if arena.cluster_shape(c)[1] >= 3][:12]
    subsets = [np.sort(rng.choice(arena.cluster_shape(c)[1], size=max(1, arena.cluster_shape(c)[1] // 2), replace=False)) for c in ids]
    parts, noise_count, total_count, _, n_classes = ctx.subset_em(arena, ids, subsets)
    for s, (c, subset) in enumerate(zip(ids, subsets)):
        ab, nc, tc, rows = reference_subset_em(reference, arena, c, subset)
        assert n_classes[s] == rows and total_count[s] == tc
        assert np.allclose(parts[s], ab, rtol=REL_TOL, atol=0) and np.array_equal(parts[s] == 0, ab == 0)
        assert abs(noise_count[s] - nc) <= REL_TOL * max(1.0, tc)


def test_subset_em_rejects_malformed_subsets(ctx):
    from rpvg_b200 import capi
    arena = synth.transcripts_arena(seed=23, n_paths=100, n_reads=500, n_max=20)
    with pytest.raises(capi.RpvgB200Error, match="malformed subsets"):
        ctx.subset_em(arena, [0], [np.array([10 ** 6], dtype=np.uint32)])
