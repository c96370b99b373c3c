"""Sharding of a cluster arena over the GPUs of one box.

Clusters are mutually independent (they are connected components, rpvg:src/path_clusters.cpp:163-207, and the reference
treats them as independent OpenMP work items, rpvg:src/main.cpp:829), so there is no collective on the data path: every
rank runs the hot path on its own clusters and the per-cluster results are gathered on the host in the original cluster
order. The assignment is longest-processing-time-first on the cost model  nnz_c x E[iterations_c]  — the reference's own
schedule is the same idea (largest first, dynamic, rpvg:src/main.cpp:827-830).
"""

from __future__ import annotations

import heapq

import numpy as np


def expected_em_iterations(weight):
    """Prior for the EM iterations of a cluster with ``weight`` = entries + read classes. Iterations grow with cluster
    size (observed on config 2: ~30 for a handful of entries, ~1000 at 1e4, 3000-5000 at 1e5); the stopping rule caps
    them at max_em_its."""
    return np.clip(12.0 * np.sqrt(np.asarray(weight, dtype=np.float64)), 20.0, 5000.0)


def cluster_costs(arena) -> np.ndarray:
    """nnz_c x E[iterations_c] per cluster (nnz = path entries + one noise entry per read class)."""
    reads = np.diff(arena.read_offset.astype(np.int64))
    bounds = arena.row_ptr[arena.read_offset.astype(np.int64)].astype(np.int64)
    weight = np.diff(bounds) + reads
    return weight * expected_em_iterations(weight)


def lpt_partition(costs, n_parts: int) -> list[np.ndarray]:
    """Longest processing time first: clusters by descending cost, each to the least loaded part. Returns, per part,
    the cluster ids in ascending order. Deterministic (ties: lower cluster id first, lower part first)."""
    costs = np.asarray(costs, dtype=np.float64)
    order = np.lexsort((np.arange(len(costs)), -costs))
    heap = [(0.0, p) for p in range(n_parts)]
    heapq.heapify(heap)
    parts = [[] for _ in range(n_parts)]
    for c in order:
        load, p = heapq.heappop(heap)
        parts[p].append(int(c))
        heapq.heappush(heap, (load + float(costs[c]), p))
    return [np.array(sorted(ids), dtype=np.int64) for ids in parts]


def run_sharded(arena, compute, rank: int, world: int, gather=None):
    """Runs ``compute(sub_arena) -> (abundances, noise_count, total_count, iterations)`` on this rank's shard and, on
    rank 0, reassembles the results of all ranks in the original cluster order (other ranks return None).

    ``gather(obj) -> list`` collects one picklable object per rank on rank 0 (``torch.distributed.gather_object``
    wrapped by the caller); with world == 1 it may be None. The exchange happens after the hot path, on the host: it is
    the reference's own "results gathered in host memory" step, not a data-path collective.
    """
    parts = lpt_partition(cluster_costs(arena), world)
    mine = parts[rank]
    local = None
    if len(mine):
        sub = arena.select(mine)
        local = compute(sub)
    payload = (mine, local)
    gathered = [payload] if world == 1 else gather(payload)
    if rank != 0:
        return None

    abundances = np.zeros(arena.n_paths, dtype=np.float64)
    noise_count = np.zeros(arena.n_clusters, dtype=np.float64)
    total_count = np.zeros(arena.n_clusters, dtype=np.float64)
    iterations = np.zeros(arena.n_clusters, dtype=np.uint32)
    path_offset = arena.path_offset.astype(np.int64)

    for ids, res in gathered:
        if res is None:
            continue
        ab, noise, total, its = res
        pos = 0
        for k, c in enumerate(ids):
            n = int(path_offset[c + 1] - path_offset[c])
            abundances[path_offset[c]:path_offset[c] + n] = ab[pos:pos + n]
            pos += n
            noise_count[c], total_count[c], iterations[c] = noise[k], total[k], its[k]
    return abundances, noise_count, total_count, iterations
