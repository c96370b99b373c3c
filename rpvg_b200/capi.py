"""ctypes view of librpvg_b200.so (include/rpvg_b200.h) for the Python test and bench drivers.

The product is the C ABI; this file only marshals numpy / pinned buffers into it. There is no fallback: if the shared
library is missing or no CUDA device is present, every entry point raises.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .arena import ClusterArena

_LIB_PATH = os.environ.get("RPVG_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "librpvg_b200.so")
_lib = None


class RpvgB200Error(RuntimeError):
    pass


class Clusters(C.Structure):
    _fields_ = [
        ("n_clusters", C.c_uint32),
        ("read_offset", C.c_void_p),
        ("path_offset", C.c_void_p),
        ("row_ptr", C.c_void_p),
        ("col", C.c_void_p),
        ("prob", C.c_void_p),
        ("noise", C.c_void_p),
        ("count", C.c_void_p),
    ]


class EmParams(C.Structure):
    _fields_ = [
        ("max_em_its", C.c_uint32),
        ("max_rel_em_conv", C.c_double),
        ("min_em_conv_its", C.c_uint32),
        ("min_em_abundance", C.c_double),
    ]


class CollapseProblems(C.Structure):
    _fields_ = [("n_problems", C.c_uint32)] + [(n, C.c_void_p) for n in ("row_offset", "row_ptr", "col", "val", "noise", "count")]


class ReadRecords(C.Structure):
    _fields_ = [("n_clusters", C.c_uint32)] + [(n, C.c_void_p) for n in ("read_offset", "group_ptr", "group_prob", "id_ptr", "ids", "noise", "count")]


class Alignments(C.Structure):
    _fields_ = [("n_clusters", C.c_uint32)] + [(n, C.c_void_p) for n in (
        "read_offset", "path_offset", "effective_length", "path_group", "source_count", "min_mapq", "align_ptr", "score_sum", "align_length",
        "frag_log_prob", "id_ptr", "ids")]


class Subsets(C.Structure):
    _fields_ = [("n_subsets", C.c_uint32), ("subset_cluster", C.c_void_p), ("subset_offset", C.c_void_p), ("subset_paths", C.c_void_p)]


def library_path() -> str:
    return _LIB_PATH


def lib():
    """Load librpvg_b200.so (built in-tree by ``__graft_entry__.build()`` / rpvg_b200/csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RpvgB200Error(f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(there is no CPU fallback for the CUDA extension)")
        l = C.CDLL(_LIB_PATH)
        l.rpvg_b200_version.restype = C.c_char_p
        l.rpvg_b200_last_error.restype = C.c_char_p
        l.rpvg_b200_last_error.argtypes = [C.c_void_p]
        l.rpvg_b200_launch_count.restype = C.c_uint64
        l.rpvg_b200_launch_count.argtypes = [C.c_void_p]
        l.rpvg_b200_pinned_scratch.restype = C.c_void_p
        l.rpvg_b200_pinned_scratch.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64]
        l.rpvg_b200_create.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
        l.rpvg_b200_destroy.argtypes = [C.c_void_p]
        l.rpvg_b200_em_params_default.argtypes = [C.POINTER(EmParams)]
        l.rpvg_b200_em_abundances.argtypes = [C.c_void_p, C.POINTER(Clusters), C.POINTER(EmParams)] + [C.c_void_p] * 4
        l.rpvg_b200_batch_upload.argtypes = [C.c_void_p, C.POINTER(Clusters), C.POINTER(C.c_void_p)]
        l.rpvg_b200_batch_em.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(EmParams)]
        l.rpvg_b200_batch_download.argtypes = [C.c_void_p, C.c_void_p] + [C.c_void_p] * 4
        l.rpvg_b200_batch_em_work.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        l.rpvg_b200_batch_em_timings.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        l.rpvg_b200_batch_free.argtypes = [C.c_void_p, C.c_void_p]
        l.rpvg_b200_merge_read_records.argtypes = [C.c_void_p, C.POINTER(ReadRecords), C.c_double] + [C.c_void_p] * 4
        l.rpvg_b200_add_path_probs.argtypes = [C.c_void_p, C.POINTER(Alignments), C.c_double, C.c_double] + [C.c_void_p] * 5
        l.rpvg_b200_tpm.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
        l.rpvg_b200_haplotype_marginals.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 8
        l.rpvg_b200_joint_haplotype_estimates.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 8 + [C.c_uint32, C.c_double, C.c_double] + [C.c_void_p] * 4
        l.rpvg_b200_collapse.argtypes = [C.c_void_p, C.POINTER(CollapseProblems), C.c_double] + [C.c_void_p] * 4
        l.rpvg_b200_subset_em.argtypes = [C.c_void_p, C.POINTER(Clusters), C.POINTER(Subsets), C.POINTER(EmParams), C.c_double] + [C.c_void_p] * 5
        l.rpvg_b200_path_cover.argtypes = [C.c_void_p, C.POINTER(Clusters), C.c_void_p]
        l.rpvg_b200_weighted_path_cover.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5
        l.rpvg_b200_em_cluster_noalloc.argtypes = ([C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint64] + [C.c_void_p] * 5 + [C.POINTER(EmParams)] +
                                                   [C.c_void_p] * 4 + [C.c_char_p, C.c_size_t])
        _lib = l
    return _lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor (pinned host memory)


def clusters_struct(arena) -> Clusters:
    """``arena`` is a ClusterArena (numpy) or any object with the same attribute names holding torch CPU tensors."""
    s = Clusters()
    s.n_clusters = len(arena.read_offset) - 1
    s.read_offset = _ptr(arena.read_offset)
    s.path_offset = _ptr(arena.path_offset)
    s.row_ptr = _ptr(arena.row_ptr)
    s.col = _ptr(arena.col)
    s.prob = _ptr(arena.prob)
    s.noise = _ptr(arena.noise)
    s.count = _ptr(arena.count)
    return s


def default_em_params(max_em_its=10000, max_rel_em_conv=0.001) -> EmParams:
    p = EmParams()
    lib().rpvg_b200_em_params_default(C.byref(p))
    p.max_em_its = max_em_its
    p.max_rel_em_conv = max_rel_em_conv
    return p


class Context:
    """One ``rpvg_b200_ctx``: a device and the stream all of its work is enqueued on."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self._h = C.c_void_p()
        status = lib().rpvg_b200_create(device, C.c_void_p(stream) if stream else None, C.byref(self._h))
        if status != 0:
            self._h = C.c_void_p()
            raise RpvgB200Error(f"rpvg_b200_create(device={device}) failed with status {status}: "
                                "a CUDA device is required (no CPU fallback)")

    def close(self):
        if self._h:
            lib().rpvg_b200_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, status: int):
        if status != 0:
            raise RpvgB200Error(f"status {status}: {lib().rpvg_b200_last_error(self._h).decode()}")

    @property
    def launches(self) -> int:
        return int(lib().rpvg_b200_launch_count(self._h))

    def pinned_scratch(self, slot: int, nbytes: int) -> int:
        """rpvg_b200_pinned_scratch: address of the slot's page-locked buffer (>= nbytes); raises on failure."""
        address = lib().rpvg_b200_pinned_scratch(self._h, slot, nbytes)
        if not address:
            raise RpvgB200Error(lib().rpvg_b200_last_error(self._h).decode())
        return int(address)

    # -- transcripts model --------------------------------------------------------------------------------------------
    def em_abundances(self, arena: ClusterArena, params: EmParams | None = None):
        """rpvg_b200_em_abundances: host in, host out. Returns (abundances, noise_count, total_count, iterations)."""
        params = params or default_em_params()
        Cn = arena.n_clusters
        abundances = np.empty(arena.n_paths, dtype=np.float64)
        noise_count = np.empty(Cn, dtype=np.float64)
        total_count = np.empty(Cn, dtype=np.float64)
        iterations = np.empty(Cn, dtype=np.uint32)
        cl = clusters_struct(arena)
        self._check(lib().rpvg_b200_em_abundances(self._h, C.byref(cl), C.byref(params), abundances.ctypes.data,
                                                  noise_count.ctypes.data, total_count.ctypes.data, iterations.ctypes.data))
        return abundances, noise_count, total_count, iterations

    def collapse(self, row_offset, row_ptr, col, val, noise, count, prob_precision=1e-8):
        """rpvg_b200_collapse on numpy arrays. Returns (row_class, n_classes, class_rep, class_count)."""
        row_offset = np.ascontiguousarray(row_offset, dtype=np.uint64)
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.uint64)
        col = np.ascontiguousarray(col, dtype=np.uint32)
        val = np.ascontiguousarray(val, dtype=np.float64)
        noise = np.ascontiguousarray(noise, dtype=np.float64)
        count = np.ascontiguousarray(count, dtype=np.float64)
        P, R = len(row_offset) - 1, int(row_offset[-1])
        row_class = np.zeros(R, dtype=np.uint32)
        n_classes = np.zeros(P, dtype=np.uint32)
        class_rep = np.zeros(max(R, 1), dtype=np.uint32)
        class_count = np.zeros(max(R, 1), dtype=np.float64)
        pr = CollapseProblems(P, row_offset.ctypes.data, row_ptr.ctypes.data, col.ctypes.data, val.ctypes.data, noise.ctypes.data, count.ctypes.data)
        self._check(lib().rpvg_b200_collapse(self._h, C.byref(pr), prob_precision, row_class.ctypes.data, n_classes.ctypes.data, class_rep.ctypes.data,
                                             class_count.ctypes.data))
        total = int(n_classes.sum())
        return row_class, n_classes, class_rep[:total], class_count[:total]

    def subset_em(self, arena, subset_cluster, subsets, params: EmParams | None = None, prob_precision=1e-8):
        """rpvg_b200_subset_em. ``subsets`` is a list of ascending path-index lists, ``subset_cluster[s]`` the cluster of subset s.
        Returns (list of abundance arrays, noise_count, total_count, iterations, n_classes)."""
        params = params or default_em_params()
        S = len(subsets)
        subset_cluster = np.ascontiguousarray(subset_cluster, dtype=np.uint32)
        subset_offset = np.zeros(S + 1, dtype=np.uint64)
        subset_offset[1:] = np.cumsum([len(x) for x in subsets])
        subset_paths = np.ascontiguousarray(np.concatenate([np.asarray(x, dtype=np.uint32) for x in subsets]), dtype=np.uint32)
        abundances = np.zeros(int(subset_offset[-1]), dtype=np.float64)
        noise_count = np.zeros(S); total_count = np.zeros(S)
        iterations = np.zeros(S, dtype=np.uint32); n_classes = np.zeros(S, dtype=np.uint32)
        cl = clusters_struct(arena)
        ss = Subsets(S, subset_cluster.ctypes.data, subset_offset.ctypes.data, subset_paths.ctypes.data)
        self._check(lib().rpvg_b200_subset_em(self._h, C.byref(cl), C.byref(ss), C.byref(params), prob_precision, abundances.ctypes.data, noise_count.ctypes.data,
                                              total_count.ctypes.data, iterations.ctypes.data, n_classes.ctypes.data))
        parts = [abundances[int(subset_offset[s]):int(subset_offset[s + 1])] for s in range(S)]
        return parts, noise_count, total_count, iterations, n_classes

    def path_cover(self, arena):
        """rpvg_b200_path_cover: boolean array [n_paths], True for the paths of the greedy weighted minimum cover."""
        in_cover = np.zeros(arena.n_paths, dtype=np.uint8)
        cl = clusters_struct(arena)
        self._check(lib().rpvg_b200_path_cover(self._h, C.byref(cl), in_cover.ctypes.data))
        return in_cover.astype(bool)

    def weighted_path_cover(self, cover, read_counts, path_weights):
        """rpvg_b200_weighted_path_cover on a dense boolean cover (rows x paths): sorted list of the picked paths."""
        cover = np.asarray(cover, dtype=bool)
        rows, n_paths = cover.shape
        row_ptr = np.zeros(rows + 1, dtype=np.uint64)
        row_ptr[1:] = np.cumsum(cover.sum(axis=1))
        col = np.ascontiguousarray(np.nonzero(cover)[1], dtype=np.uint32)
        read_counts = np.ascontiguousarray(read_counts, dtype=np.float64)
        path_weights = np.ascontiguousarray(path_weights, dtype=np.float64)
        in_cover = np.zeros(n_paths, dtype=np.uint8)
        self._check(lib().rpvg_b200_weighted_path_cover(self._h, rows, n_paths, _ptr(row_ptr), _ptr(col), _ptr(read_counts), _ptr(path_weights), _ptr(in_cover)))
        return [int(j) for j in np.nonzero(in_cover)[0]]

    def merge_read_records(self, read_offset, group_ptr, group_prob, id_ptr, ids, noise, count, prob_precision=1e-8):
        """rpvg_b200_merge_read_records: (n_merged[C], record_class[R], merged_rep[R], merged_count[R])."""
        arrays = [np.ascontiguousarray(a, dtype=t) for a, t in ((read_offset, np.uint64), (group_ptr, np.uint64), (group_prob, np.float64), (id_ptr, np.uint64),
                                                                   (ids, np.uint32), (noise, np.float64), (count, np.uint32))]
        rec = ReadRecords(len(arrays[0]) - 1, *[_ptr(a) for a in arrays])
        R = int(arrays[0][-1])
        n_merged = np.zeros(rec.n_clusters, dtype=np.uint32)
        record_class = np.zeros(R, dtype=np.uint32)
        merged_rep = np.zeros(R, dtype=np.uint32)
        merged_count = np.zeros(R, dtype=np.uint32)
        self._check(lib().rpvg_b200_merge_read_records(self._h, C.byref(rec), prob_precision, _ptr(n_merged), _ptr(record_class), _ptr(merged_rep), _ptr(merged_count)))
        return n_merged, record_class, merged_rep, merged_count

    def add_path_probs(self, read_offset, path_offset, effective_length, min_mapq, align_ptr, score_sum, align_length, frag_log_prob, id_ptr, ids,
                       path_group=None, source_count=None, min_noise_prob=1e-4, prob_precision=1e-8):
        """rpvg_b200_add_path_probs: (noise[R], group_ptr[R + 1], group_prob, group_id_ptr, group_ids)."""
        def arr(a, t):
            return None if a is None else np.ascontiguousarray(a, dtype=t)
        read_offset, path_offset, align_ptr, id_ptr = (arr(a, np.uint64) for a in (read_offset, path_offset, align_ptr, id_ptr))
        effective_length, frag_log_prob = arr(effective_length, np.float64), arr(frag_log_prob, np.float64)
        min_mapq, score_sum = arr(min_mapq, np.uint8), arr(score_sum, np.int32)
        align_length, ids, path_group, source_count = (arr(a, np.uint32) for a in (align_length, ids, path_group, source_count))
        al = Alignments(len(read_offset) - 1, _ptr(read_offset), _ptr(path_offset), _ptr(effective_length), _ptr(path_group), _ptr(source_count), _ptr(min_mapq),
                        _ptr(align_ptr), _ptr(score_sum), _ptr(align_length), _ptr(frag_log_prob), _ptr(id_ptr), _ptr(ids))
        R, T = int(read_offset[-1]), len(ids)
        noise = np.zeros(R, dtype=np.float64)
        group_ptr = np.zeros(R + 1, dtype=np.uint64)
        group_prob = np.zeros(T + 1, dtype=np.float64)
        group_id_ptr = np.zeros(T + 2, dtype=np.uint64)
        group_ids = np.zeros(T + 1, dtype=np.uint32)
        self._check(lib().rpvg_b200_add_path_probs(self._h, C.byref(al), min_noise_prob, prob_precision, _ptr(noise), _ptr(group_ptr), _ptr(group_prob),
                                                   _ptr(group_id_ptr), _ptr(group_ids)))
        G = int(group_ptr[-1])
        return noise, group_ptr, group_prob[:G], group_id_ptr[:G + 1], group_ids[:int(group_id_ptr[G])]

    def haplotype_marginals(self, path_offset, set_offset, member_offset, members, posteriors, abundances):
        """rpvg_b200_haplotype_marginals: (haplotype_probs[n_paths], read_counts[n_paths])."""
        path_offset = np.ascontiguousarray(path_offset, dtype=np.uint64)
        set_offset = np.ascontiguousarray(set_offset, dtype=np.uint64)
        member_offset = np.ascontiguousarray(member_offset, dtype=np.uint64)
        members = np.ascontiguousarray(members, dtype=np.uint32)
        posteriors = np.ascontiguousarray(posteriors, dtype=np.float64)
        abundances = np.ascontiguousarray(abundances, dtype=np.float64)
        n_paths = int(path_offset[-1])
        probs = np.zeros(n_paths, dtype=np.float64)
        counts = np.zeros(n_paths, dtype=np.float64)
        self._check(lib().rpvg_b200_haplotype_marginals(self._h, path_offset.size - 1, _ptr(path_offset), _ptr(set_offset), _ptr(member_offset), _ptr(members),
                                                        _ptr(posteriors), _ptr(abundances), _ptr(probs), _ptr(counts)))
        return probs, counts

    def joint_haplotype_estimates(self, path_offset, set_offset, member_offset, members, posteriors, abundances, effective_lengths, noise_count, ploidy,
                                  min_posterior, total_transcript_count):
        """rpvg_b200_joint_haplotype_estimates: (written[n_sets] bool, read_counts[n_members], tpm[n_members], noise_counts[ploidy])."""
        path_offset, set_offset, member_offset = (np.ascontiguousarray(a, dtype=np.uint64) for a in (path_offset, set_offset, member_offset))
        members = np.ascontiguousarray(members, dtype=np.uint32)
        posteriors, abundances, effective_lengths, noise_count = (np.ascontiguousarray(a, dtype=np.float64) for a in (posteriors, abundances, effective_lengths, noise_count))
        S, E = int(set_offset[-1]), int(member_offset[-1])
        written = np.zeros(max(S, 1), dtype=np.uint8)
        read_counts = np.zeros(max(E, 1), dtype=np.float64)
        tpm = np.zeros(max(E, 1), dtype=np.float64)
        noise_counts = np.zeros(ploidy, dtype=np.float64)
        self._check(lib().rpvg_b200_joint_haplotype_estimates(self._h, path_offset.size - 1, _ptr(path_offset), _ptr(set_offset), _ptr(member_offset), _ptr(members),
                                                              _ptr(posteriors), _ptr(abundances), _ptr(effective_lengths), _ptr(noise_count), ploidy, min_posterior,
                                                              total_transcript_count, _ptr(written), _ptr(read_counts), _ptr(tpm), _ptr(noise_counts)))
        return written[:S].astype(bool), read_counts[:E], tpm[:E], noise_counts

    def tpm(self, abundances, effective_lengths):
        """rpvg_b200_tpm: (tpm[n_paths], total_transcript_count)."""
        abundances = np.ascontiguousarray(abundances, dtype=np.float64)
        effective_lengths = np.ascontiguousarray(effective_lengths, dtype=np.float64)
        assert abundances.shape == effective_lengths.shape
        out = np.zeros(abundances.size, dtype=np.float64)
        total = C.c_double(0)
        self._check(lib().rpvg_b200_tpm(self._h, abundances.size, _ptr(abundances), _ptr(effective_lengths), _ptr(out), C.byref(total)))
        return out, total.value

    def upload(self, arena) -> "Batch":
        cl = clusters_struct(arena)
        h = C.c_void_p()
        self._check(lib().rpvg_b200_batch_upload(self._h, C.byref(cl), C.byref(h)))
        return Batch(self, h)


class Batch:
    """A cluster arena resident in HBM (``rpvg_b200_batch``)."""

    def __init__(self, ctx: Context, handle):
        self.ctx = ctx
        self._h = handle

    def em(self, params: EmParams | None = None):
        params = params or default_em_params()
        self.ctx._check(lib().rpvg_b200_batch_em(self.ctx._h, self._h, C.byref(params)))

    def download(self, abundances, noise_count, total_count, iterations):
        self.ctx._check(lib().rpvg_b200_batch_download(self.ctx._h, self._h, _ptr(abundances), _ptr(noise_count),
                                                       _ptr(total_count), _ptr(iterations)))

    def em_work(self) -> tuple[float, float]:
        """(sum nnz_c x iterations_c, algorithmic bytes) of the last em()."""
        work, nbytes = C.c_double(), C.c_double()
        self.ctx._check(lib().rpvg_b200_batch_em_work(self.ctx._h, self._h, C.byref(work), C.byref(nbytes)))
        return work.value, nbytes.value

    def em_timings(self) -> tuple[float, float]:
        """(build_ms, em_ms): device time of the matrix build and of the EM kernels of the last em()."""
        build, em = C.c_float(), C.c_float()
        self.ctx._check(lib().rpvg_b200_batch_em_timings(self.ctx._h, self._h, C.byref(build), C.byref(em)))
        return build.value, em.value

    def free(self):
        if self._h:
            lib().rpvg_b200_batch_free(self.ctx._h, self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def em_cluster_noalloc(device, stream, n_reads, n_paths, nnz, d_row_ptr, d_col, d_prob, d_noise, d_count, params, d_abundances, d_noise_count,
                       d_total_count, d_iterations):
    """rpvg_b200_em_cluster_noalloc: every pointer argument is a device address (int) owned by the caller; nothing is
    allocated. Returns (status, message)."""
    message = C.create_string_buffer(512)
    status = lib().rpvg_b200_em_cluster_noalloc(device, C.c_void_p(stream) if stream else None, n_reads, n_paths, nnz, d_row_ptr, d_col, d_prob, d_noise,
                                                d_count, C.byref(params), d_abundances, d_noise_count, d_total_count, d_iterations, message, len(message))
    return status, message.value.decode()
