"""Frozen reference output (tests/golden) against the CPU port (CPU test) and the CUDA path (GPU test)."""
import numpy as np
import pytest

from oracle import loader
from tests.golden_util import cluster_results, load


def check_transcripts(arena, z, abundances, noise_count, total_count, tol):
    for c in range(arena.n_clusters):
        _, post, ref_ab, ref_noise, ref_total = cluster_results(z, "", c)
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        mine = abundances[p0:p1]
        assert np.all(post == 0)
        assert np.array_equal(mine == 0, ref_ab == 0), f"cluster {c}: set of paths above min_em_abundance differs"
        nz = ref_ab != 0
        if nz.any():
            assert np.max(np.abs(mine[nz] - ref_ab[nz]) / ref_ab[nz]) < tol
        assert abs(noise_count[c] - ref_noise) <= tol * max(1.0, ref_total)
        assert total_count[c] == ref_total


def test_port_reproduces_golden_transcripts_em():
    arena, z = load("transcripts_em.npz")
    ab, noise, total, _, _ = loader.port_em_abundances(arena)
    check_transcripts(arena, z, ab, noise, total, 1e-10)


def test_golden_haplotype_models_are_consistent():
    """Sanity of the frozen haplotype-model fixtures: posteriors sum to <= 1, read counts are conserved."""
    arena, z = load("haplotype_models.npz")
    for c in range(arena.n_clusters):
        sets, post, _, _, _ = cluster_results(z, "haplotypes_", c)
        assert len(sets) == len(post) and abs(post.sum() - 1) < 1e-9
        sets, post, ab, noise, total = cluster_results(z, "nested_", c)
        assert abs(ab.sum() + noise - total) <= 1e-6 * total
        sets, post, ab, noise, total = cluster_results(z, "strains_", c)
        assert abs(ab.sum() + noise - total) <= 1e-6 * max(1.0, total)


@pytest.mark.gpu
def test_cuda_em_reproduces_golden_transcripts_em(ctx):
    arena, z = load("transcripts_em.npz")
    ab, noise, total, _ = ctx.em_abundances(arena)
    check_transcripts(arena, z, ab, noise, total, 1e-6)
