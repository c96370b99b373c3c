"""Generates tests/golden/*.npz from the reference's own compiled code (oracle/_ref). Run where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden.py

Each fixture = a small seeded cluster arena + what the reference's estimator classes return for it, so that the GPU box
(which has no reference sources) can check the CUDA path and the CPU port against frozen reference output.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader  # noqa: E402
from rpvg_b200 import synth  # noqa: E402
from rpvg_b200.arena import concat  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ARENA_FIELDS = ("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count", "group_id", "source_count", "source_ptr", "source_ids")


def arena_dict(arena):
    return {f"arena_{k}": getattr(arena, k) for k in ARENA_FIELDS if getattr(arena, k) is not None}


def flatten_results(results, prefix=""):
    out = {}
    out[prefix + "set_ptr"] = np.cumsum([0] + [len(r["sets"]) for r in results]).astype(np.uint64)
    sizes = [len(s) for r in results for s in r["sets"]]
    out[prefix + "set_size"] = np.array(sizes, dtype=np.uint32)
    out[prefix + "set_ids"] = np.array([x for r in results for s in r["sets"] for x in s], dtype=np.uint32)
    out[prefix + "post_ptr"] = np.cumsum([0] + [len(r["posteriors"]) for r in results]).astype(np.uint64)
    out[prefix + "posteriors"] = np.concatenate([r["posteriors"] for r in results]) if results else np.zeros(0)
    out[prefix + "abund_ptr"] = np.cumsum([0] + [len(r["abundances"]) for r in results]).astype(np.uint64)
    out[prefix + "abundances"] = np.concatenate([r["abundances"] for r in results]) if results else np.zeros(0)
    out[prefix + "noise_count"] = np.array([r["noise_count"] for r in results])
    out[prefix + "total_count"] = np.array([r["total_count"] for r in results])
    return out


def main():
    ref = loader.reference()

    # 1. transcripts model: EM on power-law clusters
    arena = synth.transcripts_arena(seed=101, n_paths=600, n_reads=3000, n_max=120)
    _, res = loader.harness_estimate_batch(ref, loader.make_config(loader.MODEL_TRANSCRIPTS), arena)
    np.savez_compressed(os.path.join(HERE, "transcripts_em.npz"), **arena_dict(arena), **flatten_results(res))

    # 2. haplotype clusters (config-1 style): haplotypes model (bounded posteriors), haplotype-transcripts (nested), strains
    rng = np.random.default_rng(202)
    hap = concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 4)), paths_per_transcript=(2, 7),
                                          n_haplotypes=24, n_reads=int(rng.integers(60, 300))) for _ in range(12)])
    out = arena_dict(hap)
    for name, cfg in [
        ("haplotypes", loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2)),
        ("nested", loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)),
        ("strains", loader.make_config(loader.MODEL_STRAINS)),
        ("path_posterior", loader.make_config(loader.MODEL_PATH_POSTERIOR)),
    ]:
        _, res = loader.harness_estimate_batch(ref, cfg, hap, rng_seed=7)
        out.update(flatten_results(res, prefix=name + "_"))
    np.savez_compressed(os.path.join(HERE, "haplotype_models.npz"), **out)

    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
