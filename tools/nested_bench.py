"""Inference-stage wall time of the haplotype-transcripts model (config-1-like synthetic clusters) through the SAME harness
source compiled against the reference's classes (CPU, oracle/_ref) and against this repository's drop-in classes (GPU).
python tools/nested_bench.py [n_clusters] [threads]
RPVG_BENCH_GPU_THREADS="16,64,256": caller threads of the per-cluster estimate() loop on the GPU classes (they block while
the device works, so more of them than cores is how a per-cluster interface keeps the device fed)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import loader
from rpvg_b200 import synth
from rpvg_b200.arena import concat

n_clusters = int(sys.argv[1]) if len(sys.argv) > 1 else 400
threads = int(sys.argv[2]) if len(sys.argv) > 2 else 16
# the generator is slow Python: the arena is cached under tools/cache (git-ignored, travels to the GPU box)
cache = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cache", f"nested_{n_clusters}.npz")
t0 = time.time()
if os.path.exists(cache):
    z = np.load(cache)
    from rpvg_b200.arena import ClusterArena
    arena = ClusterArena(z["read_offset"], z["path_offset"], z["row_ptr"], z["col"], z["prob"], z["noise"], z["count"], group_id=z["group_id"],
                         source_count=z["source_count"], source_ptr=z["source_ptr"], source_ids=z["source_ids"])
else:
    rng = np.random.default_rng(20240601)
    arena = concat([synth.haplotype_cluster(rng, n_transcripts=int(rng.integers(1, 5)), paths_per_transcript=(2, 24), n_haplotypes=60,
                                            n_reads=int(rng.integers(200, 3000))) for _ in range(n_clusters)])
    os.makedirs(os.path.dirname(cache), exist_ok=True)
    np.savez_compressed(cache, read_offset=arena.read_offset, path_offset=arena.path_offset, row_ptr=arena.row_ptr, col=arena.col, prob=arena.prob, noise=arena.noise,
                        count=arena.count, group_id=arena.group_id, source_count=arena.source_count, source_ptr=arena.source_ptr, source_ids=arena.source_ids)
print(f"{n_clusters} clusters, {arena.n_paths} paths, {arena.n_reads} read classes, ready in {time.time() - t0:.1f} s", flush=True)
if len(sys.argv) > 3 and sys.argv[3] == "generate-only":
    raise SystemExit(0)
model = os.environ.get("RPVG_BENCH_MODEL", "haplotype-transcripts")
cfg = {"haplotype-transcripts": loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2),
       "haplotypes": loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2),
       "strains": loader.make_config(loader.MODEL_STRAINS),
       "transcripts": loader.make_config(loader.MODEL_TRANSCRIPTS)}[model]
print("inference model:", model, flush=True)

b200 = loader.load_harness(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "_build", "libharness_b200.so"))
batch_only = len(sys.argv) > 3 and sys.argv[3] == "batch-only"
results = None
for name, lib_ in (("reference classes (CPU)", loader.reference()), ("rpvg_b200 classes (GPU)", b200)):
    if lib_ is None or batch_only:
        continue
    counts = [threads]
    if lib_ is b200 and os.environ.get("RPVG_BENCH_GPU_THREADS"):
        counts = [int(x) for x in os.environ["RPVG_BENCH_GPU_THREADS"].split(",")]
    for n_threads in counts:
        best = None
        for rep in range(2):
            secs, results = loader.harness_estimate_batch(lib_, cfg, arena, rng_seed=1, num_threads=n_threads)
            best = secs if best is None else min(best, secs)
        n_sets = sum(len(r["sets"]) for r in results)
        print(f"{name}: {best * 1e3:.1f} ms for {n_clusters} clusters with {n_threads} threads = {n_clusters / best:.0f} clusters/s ({n_sets} output tuples)", flush=True)

best = None
for rep in range(3):
    secs, batched = loader.harness_estimate_batched(b200, cfg, arena, rng_seed=1)
    best = secs if best is None else min(best, secs)
print(f"rpvg_b200 classes, estimateBatch (GPU, one call): {best * 1e3:.1f} ms = {n_clusters / best:.0f} clusters/s", flush=True)
if results is not None:
    same = all(a["sets"] == b["sets"] and np.array_equal(a["abundances"], b["abundances"]) for a, b in zip(batched, results))
    print("batched results identical to the per-cluster ones:", same)
