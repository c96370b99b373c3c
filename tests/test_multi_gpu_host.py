"""estimateBatchOnDevices: the C++ host library deals a batch over the GPUs of the box (LPT on a per-cluster cost, one host
thread + context per device, results written into the caller's objects). Needs at least 2 devices; the results must be
those of the single-device call bit for bit, in the caller's order."""

import numpy as np
import pytest

from tests.harness import binding
from rpvg_b200 import synth
from rpvg_b200.arena import concat

pytestmark = pytest.mark.gpu


def device_count():
    from rpvg_b200 import capi
    lib = capi.lib()
    return int(lib.rpvg_b200_device_count())


def same(a, b):
    assert a["sets"] == b["sets"]
    assert np.array_equal(a["posteriors"], b["posteriors"])
    assert np.array_equal(a["abundances"], b["abundances"])
    assert a["noise_count"] == b["noise_count"] and a["total_count"] == b["total_count"]


@pytest.mark.parametrize("model", ["transcripts", "strains", "haplotypes", "haplotype-transcripts"])
def test_batch_over_two_devices_is_identical(b200_harness, model):
    if device_count() < 2:
        pytest.skip("needs two GPUs")

    if model in ("transcripts", "strains"):
        arena = synth.transcripts_arena(seed=41, n_paths=6000, n_reads=30000, n_max=400)
        cfg = binding.make_config(binding.MODEL_TRANSCRIPTS if model == "transcripts" else binding.MODEL_STRAINS)
    else:
        rng = np.random.default_rng(17)
        arena = concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 4)), paths_per_transcript=(2, 8), n_haplotypes=30,
                                                n_reads=int(rng.integers(100, 500))) for _ in range(40)])
        cfg = binding.make_config(binding.MODEL_HAPLOTYPES if model == "haplotypes" else binding.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)

    _, one = binding.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=3, num_devices=1)
    _, two = binding.harness_estimate_batched(b200_harness, cfg, arena, rng_seed=3, num_devices=2)
    assert len(one) == len(two) == arena.n_clusters
    for a, b in zip(one, two):
        same(a, b)


def test_device_count_entry_point():
    assert device_count() >= 1
