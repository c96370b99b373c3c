"""Cluster arena: the flattened input of the hot path (include/rpvg_b200.h, ``rpvg_b200_clusters``).

A batch of path clusters as one CSR over read classes. This is rpvg's ``vector<ReadPathProbabilities>``
(rpvg:src/read_path_probabilities.hpp:19-44) per cluster with every ``pair<double, vector<uint32_t>>`` expanded to
one entry per path id, entries sorted by path index inside a read class.
"""

from __future__ import annotations

import dataclasses

import numpy as np


@dataclasses.dataclass
class ClusterArena:

    read_offset: np.ndarray  # u64 [C + 1]
    path_offset: np.ndarray  # u64 [C + 1]
    row_ptr: np.ndarray      # u64 [R + 1]
    col: np.ndarray          # u32 [nnz]
    prob: np.ndarray         # f64 [nnz]
    noise: np.ndarray        # f64 [R]
    count: np.ndarray        # u32 [R]
    # per-path metadata (PathInfo, rpvg:src/path_cluster_estimates.hpp:14-32); optional
    group_id: np.ndarray | None = None       # u32 [P] transcript id
    source_count: np.ndarray | None = None   # u32 [P]
    source_ptr: np.ndarray | None = None     # u64 [P + 1]
    source_ids: np.ndarray | None = None     # u32 haplotype ids

    def __post_init__(self):
        self.read_offset = np.ascontiguousarray(self.read_offset, dtype=np.uint64)
        self.path_offset = np.ascontiguousarray(self.path_offset, dtype=np.uint64)
        self.row_ptr = np.ascontiguousarray(self.row_ptr, dtype=np.uint64)
        self.col = np.ascontiguousarray(self.col, dtype=np.uint32)
        self.prob = np.ascontiguousarray(self.prob, dtype=np.float64)
        self.noise = np.ascontiguousarray(self.noise, dtype=np.float64)
        self.count = np.ascontiguousarray(self.count, dtype=np.uint32)

    @property
    def n_clusters(self) -> int:
        return len(self.read_offset) - 1

    @property
    def n_reads(self) -> int:
        return int(self.read_offset[-1])

    @property
    def n_paths(self) -> int:
        return int(self.path_offset[-1])

    @property
    def nnz_path(self) -> int:
        return int(self.row_ptr[-1])

    @property
    def nnz(self) -> int:
        """Structural non-zeros including one noise entry per read class (SURVEY.md §8 notation)."""
        return self.nnz_path + self.n_reads

    def input_bytes(self) -> int:
        arrays = [self.read_offset, self.path_offset, self.row_ptr, self.col, self.prob, self.noise, self.count]
        return int(sum(a.nbytes for a in arrays))

    def cluster_shape(self, c: int) -> tuple[int, int, int]:
        """(M read classes, N paths, path non-zeros) of cluster c."""
        r0, r1 = int(self.read_offset[c]), int(self.read_offset[c + 1])
        return r1 - r0, int(self.path_offset[c + 1] - self.path_offset[c]), int(self.row_ptr[r1] - self.row_ptr[r0])

    def cluster(self, c: int) -> "ClusterArena":
        """Cluster c as an arena of its own (copies)."""
        return self.select([c])

    def select(self, ids) -> "ClusterArena":
        """The clusters in ``ids`` (in that order) as a new arena."""
        read_offset, path_offset, row_ptr = [0], [0], [np.zeros(1, dtype=np.uint64)]
        col, prob, noise, count = [], [], [], []
        group_id, source_count, source_ptr, source_ids = [], [], [np.zeros(1, dtype=np.uint64)], []
        nnz_base = 0
        src_base = 0
        for c in ids:
            r0, r1 = int(self.read_offset[c]), int(self.read_offset[c + 1])
            p0, p1 = int(self.path_offset[c]), int(self.path_offset[c + 1])
            e0, e1 = int(self.row_ptr[r0]), int(self.row_ptr[r1])
            read_offset.append(read_offset[-1] + (r1 - r0))
            path_offset.append(path_offset[-1] + (p1 - p0))
            row_ptr.append(self.row_ptr[r0 + 1:r1 + 1] - np.uint64(e0) + np.uint64(nnz_base))
            nnz_base += e1 - e0
            col.append(self.col[e0:e1])
            prob.append(self.prob[e0:e1])
            noise.append(self.noise[r0:r1])
            count.append(self.count[r0:r1])
            if self.group_id is not None:
                group_id.append(self.group_id[p0:p1])
            if self.source_count is not None:
                source_count.append(self.source_count[p0:p1])
            if self.source_ptr is not None:
                s0, s1 = int(self.source_ptr[p0]), int(self.source_ptr[p1])
                source_ptr.append(self.source_ptr[p0 + 1:p1 + 1] - np.uint64(s0) + np.uint64(src_base))
                src_base += s1 - s0
                source_ids.append(self.source_ids[s0:s1])

        def cat(parts, dtype):
            return np.concatenate(parts).astype(dtype) if parts else np.zeros(0, dtype=dtype)

        return ClusterArena(
            read_offset=np.array(read_offset, dtype=np.uint64), path_offset=np.array(path_offset, dtype=np.uint64),
            row_ptr=cat(row_ptr, np.uint64), col=cat(col, np.uint32), prob=cat(prob, np.float64), noise=cat(noise, np.float64),
            count=cat(count, np.uint32),
            group_id=cat(group_id, np.uint32) if self.group_id is not None else None,
            source_count=cat(source_count, np.uint32) if self.source_count is not None else None,
            source_ptr=cat(source_ptr, np.uint64) if self.source_ptr is not None else None,
            source_ids=cat(source_ids, np.uint32) if self.source_ptr is not None else None)

    def dense(self, c: int) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
        """Cluster c as constructProbabilityMatrix leaves it (rpvg:src/path_estimator.cpp:55-77):
        (M x N matrix, noise [M], counts [M] as float64)."""
        M, N, _ = self.cluster_shape(c)
        r0 = int(self.read_offset[c])
        mat = np.zeros((M, N), dtype=np.float64)
        for i in range(M):
            e0, e1 = int(self.row_ptr[r0 + i]), int(self.row_ptr[r0 + i + 1])
            mat[i, self.col[e0:e1]] = self.prob[e0:e1]
        return mat, self.noise[r0:r0 + M].copy(), self.count[r0:r0 + M].astype(np.float64)


def concat(arenas) -> ClusterArena:
    """Concatenate arenas (cluster order preserved)."""
    arenas = list(arenas)
    read_offset, path_offset, row_ptr = [np.zeros(1, np.uint64)], [np.zeros(1, np.uint64)], [np.zeros(1, np.uint64)]
    r_base = p_base = e_base = 0
    for a in arenas:
        read_offset.append(a.read_offset[1:] + np.uint64(r_base))
        path_offset.append(a.path_offset[1:] + np.uint64(p_base))
        row_ptr.append(a.row_ptr[1:] + np.uint64(e_base))
        r_base += a.n_reads
        p_base += a.n_paths
        e_base += a.nnz_path
    meta = all(a.group_id is not None and a.source_ptr is not None for a in arenas)
    source_ptr = None
    if meta:
        parts, s_base = [np.zeros(1, np.uint64)], 0
        for a in arenas:
            parts.append(a.source_ptr[1:] + np.uint64(s_base))
            s_base += int(a.source_ptr[-1])
        source_ptr = np.concatenate(parts)
    return ClusterArena(
        read_offset=np.concatenate(read_offset), path_offset=np.concatenate(path_offset), row_ptr=np.concatenate(row_ptr),
        col=np.concatenate([a.col for a in arenas]), prob=np.concatenate([a.prob for a in arenas]),
        noise=np.concatenate([a.noise for a in arenas]), count=np.concatenate([a.count for a in arenas]),
        group_id=np.concatenate([a.group_id for a in arenas]) if meta else None,
        source_count=np.concatenate([a.source_count for a in arenas]) if meta else None,
        source_ptr=source_ptr,
        source_ids=np.concatenate([a.source_ids for a in arenas]) if meta else None)
