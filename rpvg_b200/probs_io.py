"""ctypes view of the `_probs.txt.gz` reader/writer in librpvg_b200_host.so (rpvg_b200/host/probs_io.hpp): rpvg's
`--write-probs` dump <-> ClusterArena + path table. Test and tool use only."""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .arena import ClusterArena

_HOST_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "librpvg_b200_host.so")
_lib = None


class _View(C.Structure):
    _fields_ = [("n_clusters", C.c_uint32), ("n_reads", C.c_uint64), ("n_paths", C.c_uint64), ("nnz", C.c_uint64)] + \
               [(n, C.c_void_p) for n in ("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count", "path_length", "path_effective_length")] + \
               [("path_names", C.c_char_p)]


def _host():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_HOST_LIB)
        _lib.rpvg_b200_probs_read.restype = C.c_void_p
        _lib.rpvg_b200_probs_read.argtypes = [C.c_char_p, C.POINTER(_View), C.c_char_p, C.c_uint64]
        _lib.rpvg_b200_probs_free.argtypes = [C.c_void_p]
        _lib.rpvg_b200_probs_write.argtypes = [C.c_char_p, C.POINTER(_View), C.c_double, C.c_char_p, C.c_uint64]
    return _lib


def _copy(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dtype))), shape=(n,)).copy()


def read_probs(filename: str):
    """-> (ClusterArena, names, lengths, effective_lengths)."""
    view = _View()
    err = C.create_string_buffer(512)
    handle = _host().rpvg_b200_probs_read(os.fsencode(filename), C.byref(view), err, 512)
    if not handle:
        raise ValueError(err.value.decode())
    try:
        nc, nr, npth, nnz = view.n_clusters, view.n_reads, view.n_paths, view.nnz
        arena = ClusterArena(_copy(view.read_offset, nc + 1, np.uint64), _copy(view.path_offset, nc + 1, np.uint64), _copy(view.row_ptr, nr + 1, np.uint64),
                             _copy(view.col, nnz, np.uint32), _copy(view.prob, nnz, np.float64), _copy(view.noise, nr, np.float64), _copy(view.count, nr, np.uint32))
        names = view.path_names.decode().split("\n") if npth else []
        return arena, names, _copy(view.path_length, npth, np.uint32), _copy(view.path_effective_length, npth, np.float64)
    finally:
        _host().rpvg_b200_probs_free(handle)


def write_probs(filename: str, arena: ClusterArena, names, lengths, effective_lengths, prob_precision: float = 1e-8):
    lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
    effective_lengths = np.ascontiguousarray(effective_lengths, dtype=np.float64)
    keep = [np.ascontiguousarray(a) for a in (arena.read_offset, arena.path_offset, arena.row_ptr, arena.col, arena.prob, arena.noise, arena.count)]
    view = _View()
    view.n_clusters, view.n_reads, view.n_paths, view.nnz = arena.n_clusters, arena.n_reads, arena.n_paths, arena.nnz_path
    for name, a in zip(("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count"), keep):
        setattr(view, name, a.ctypes.data)
    view.path_length = lengths.ctypes.data
    view.path_effective_length = effective_lengths.ctypes.data
    view.path_names = "\n".join(names).encode()
    err = C.create_string_buffer(512)
    if _host().rpvg_b200_probs_write(os.fsencode(filename), C.byref(view), prob_precision, err, 512) != 0:
        raise ValueError(err.value.decode())
