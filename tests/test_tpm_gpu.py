"""TPM reducer (rpvg_b200_tpm) against a numpy restatement of the reference's arithmetic
(rpvg:src/main.cpp:1029-1057 and rpvg:src/threaded_output_writer.cpp:305-320). north_star tolerance: 1e-6 relative."""

import numpy as np
import pytest

from rpvg_b200 import synth

pytestmark = pytest.mark.gpu


def reference_tpm(abundances, effective_lengths):
    total = 0.0
    for a, l in zip(abundances, effective_lengths):   # the reference's sequential sum
        if l > 0:
            total += a / l
    counts = np.where(effective_lengths > 0, abundances / np.where(effective_lengths > 0, effective_lengths, 1.0), 0.0)
    return counts / total * 10 ** 6, total


@pytest.mark.parametrize("n", [1, 7, 1000, 200_000])
def test_tpm_matches_reference_arithmetic(ctx, n):
    rng = np.random.default_rng(n)
    abundances = rng.gamma(0.3, 50.0, size=n)
    abundances[rng.random(n) < 0.4] = 0.0
    lengths = rng.uniform(50.0, 5000.0, size=n)
    lengths[rng.random(n) < 0.05] = 0.0          # paths shorter than the fragment length: no effective length
    if not np.any((lengths > 0) & (abundances > 0)):
        abundances[0], lengths[0] = 3.0, 100.0
    tpm, total = ctx.tpm(abundances, lengths)
    ref, ref_total = reference_tpm(abundances, lengths)
    assert abs(total - ref_total) <= 1e-12 * ref_total
    assert np.array_equal(tpm == 0, ref == 0)
    nz = ref != 0
    assert np.max(np.abs(tpm[nz] - ref[nz]) / ref[nz]) < 1e-12
    assert abs(tpm.sum() - 1e6) < 1e-3


def test_tpm_of_em_abundances(ctx):
    """End of the transcripts pipeline: EM read counts -> TPM, per-path values within 1e-6 of the CPU oracle's."""
    from oracle import loader
    arena = synth.transcripts_arena(seed=11, n_paths=2000, n_reads=10000, n_max=200)
    abundances, _, _, _ = ctx.em_abundances(arena)
    o_ab, _, _, _, _ = loader.port_em_abundances(arena)
    lengths = np.random.default_rng(5).uniform(100.0, 3000.0, size=arena.n_paths)
    tpm, _ = ctx.tpm(abundances, lengths)
    ref, _ = reference_tpm(o_ab, lengths)
    nz = ref != 0
    assert np.array_equal(tpm == 0, ref == 0)
    assert np.max(np.abs(tpm[nz] - ref[nz]) / ref[nz]) < 1e-6
