"""ctypes binding of the estimator harness ABI (tests/harness/estimator_harness.h): one C++ source compiled against this
repository's drop-in classes (tests/_build/libharness_b200.so) and, by the oracle build, against the reference's own
classes. Test and bench infrastructure: marshals a ClusterArena into the arguments of PathEstimator::estimate and flattens
the PathClusterEstimates that come back. Nothing here knows which of the two libraries it talks to."""

from __future__ import annotations

import ctypes as C

import numpy as np


class HarnessCluster(C.Structure):
    _fields_ = [("num_paths", C.c_uint32), ("num_reads", C.c_uint32)] + [(n, C.c_void_p) for n in (
        "read_count", "noise_prob", "entry_ptr", "entry_prob", "entry_path_ptr", "entry_path_ids",
        "path_group_id", "path_source_count", "path_source_ptr", "path_source_ids")]


class HarnessConfig(C.Structure):
    _fields_ = [("model", C.c_int32), ("group_size", C.c_uint32), ("min_hap_prob", C.c_double), ("infer_collapsed", C.c_int32),
                ("use_group_post_gibbs", C.c_int32), ("max_em_its", C.c_uint32), ("max_rel_em_conv", C.c_double),
                ("num_gibbs_samples", C.c_uint32), ("gibbs_thin_its", C.c_uint32), ("prob_precision", C.c_double)]


class HarnessResult(C.Structure):
    _fields_ = [("n_sets", C.c_uint64), ("set_ptr", C.POINTER(C.c_uint64)), ("set_ids", C.POINTER(C.c_uint32)),
                ("n_posteriors", C.c_uint64), ("posteriors", C.POINTER(C.c_double)),
                ("n_abundances", C.c_uint64), ("abundances", C.POINTER(C.c_double)),
                ("noise_count", C.c_double), ("total_count", C.c_double),
                ("n_gibbs", C.c_uint64), ("gibbs_path_ptr", C.POINTER(C.c_uint64)), ("gibbs_path_ids", C.POINTER(C.c_uint32)),
                ("gibbs_noise_ptr", C.POINTER(C.c_uint64)), ("gibbs_noise", C.POINTER(C.c_double)),
                ("gibbs_abund_ptr", C.POINTER(C.c_uint64)), ("gibbs_abund", C.POINTER(C.c_double))]



MODEL_TRANSCRIPTS, MODEL_STRAINS, MODEL_HAPLOTYPE_TRANSCRIPTS, MODEL_PATH_POSTERIOR, MODEL_HAPLOTYPES = range(5)


def make_config(model, group_size=2, min_hap_prob=0.001, infer_collapsed=True, use_group_post_gibbs=False, max_em_its=10000,
                max_rel_em_conv=0.001, num_gibbs_samples=0, gibbs_thin_its=25, prob_precision=1e-8) -> HarnessConfig:
    """Defaults are rpvg's command-line defaults (rpvg:src/main.cpp:402-418)."""
    return HarnessConfig(model, group_size, min_hap_prob, int(infer_collapsed), int(use_group_post_gibbs), max_em_its,
                         max_rel_em_conv, num_gibbs_samples, gibbs_thin_its, prob_precision)


def _bind_harness(l):
    l.harness_impl_name.restype = C.c_char_p
    l.harness_max_threads.restype = C.c_int32
    l.harness_set_threads.argtypes = [C.c_int32]
    l.harness_set_threads.restype = None
    l.harness_source_groups.argtypes = [C.c_uint32] + [C.c_void_p] * 5
    l.harness_source_groups.restype = C.c_uint32
    l.harness_estimator_create.restype = C.c_void_p
    l.harness_estimator_create.argtypes = [C.POINTER(HarnessConfig)]
    l.harness_estimator_destroy.argtypes = [C.c_void_p]
    l.harness_estimate.restype = C.c_void_p
    l.harness_estimate.argtypes = [C.c_void_p, C.POINTER(HarnessCluster), C.c_uint32]
    l.harness_estimate_batch.restype = C.c_double
    l.harness_estimate_batch.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(HarnessCluster), C.c_uint32, C.c_int32, C.POINTER(C.c_void_p)]
    l.harness_result_view.restype = C.POINTER(HarnessResult)
    l.harness_result_view.argtypes = [C.c_void_p]
    l.harness_result_free.argtypes = [C.c_void_p]
    return l


def load_harness(path: str):
    return _bind_harness(C.CDLL(path))


class HarnessClusterData:
    """Cluster ``c`` of an arena as the reference's own input structures: per read class a list of
    (probability, [path ids]) pairs sorted ascending, equal probabilities sharing one pair — what
    ReadPathProbabilities::addPathProbs leaves behind (rpvg:src/read_path_probabilities.cpp:184-219)."""

    def __init__(self, arena, c: int):
        r0, r1 = int(arena.read_offset[c]), int(arena.read_offset[c + 1])
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        M = r1 - r0
        entry_ptr = [0]
        entry_prob = []
        entry_path_ptr = [0]
        entry_path_ids = []
        for r in range(r0, r1):
            e0, e1 = int(arena.row_ptr[r]), int(arena.row_ptr[r + 1])
            cols = arena.col[e0:e1]
            probs = arena.prob[e0:e1]
            order = np.lexsort((cols, probs))
            prev = None
            for k in order:
                if prev is None or probs[k] != prev:
                    if prev is not None:
                        entry_path_ptr.append(len(entry_path_ids))
                    entry_prob.append(probs[k])
                    prev = probs[k]
                entry_path_ids.append(cols[k])
            if prev is not None:
                entry_path_ptr.append(len(entry_path_ids))
            entry_ptr.append(len(entry_prob))
        self.read_count = np.ascontiguousarray(arena.count[r0:r1], dtype=np.uint32)
        self.noise_prob = np.ascontiguousarray(arena.noise[r0:r1], dtype=np.float64)
        self.entry_ptr = np.array(entry_ptr, dtype=np.uint64)
        self.entry_prob = np.array(entry_prob, dtype=np.float64)
        self.entry_path_ptr = np.array(entry_path_ptr, dtype=np.uint64)
        self.entry_path_ids = np.array(entry_path_ids, dtype=np.uint32)
        self.group_id = None if arena.group_id is None else np.ascontiguousarray(arena.group_id[p0:p1], dtype=np.uint32)
        self.source_count = None if arena.source_count is None else np.ascontiguousarray(arena.source_count[p0:p1], dtype=np.uint32)
        if arena.source_ptr is not None:
            s0, s1 = int(arena.source_ptr[p0]), int(arena.source_ptr[p1])
            self.source_ptr = np.ascontiguousarray(arena.source_ptr[p0:p1 + 1] - np.uint64(s0), dtype=np.uint64)
            self.source_ids = np.ascontiguousarray(arena.source_ids[s0:s1], dtype=np.uint32)
        else:
            self.source_ptr = None
            self.source_ids = None

        def p(a):
            return None if a is None else a.ctypes.data

        self.struct = HarnessCluster(p1 - p0, M, p(self.read_count), p(self.noise_prob), p(self.entry_ptr), p(self.entry_prob),
                                     p(self.entry_path_ptr), p(self.entry_path_ids), p(self.group_id), p(self.source_count),
                                     p(self.source_ptr), p(self.source_ids))


def result_to_dict(lib_, handle, free=True) -> dict:
    v = lib_.harness_result_view(handle).contents

    def arr(ptr, n, dtype):
        return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True) if n else np.zeros(0, dtype=dtype)

    set_ptr = arr(v.set_ptr, v.n_sets + 1, np.uint64)
    set_ids = arr(v.set_ids, int(set_ptr[-1]), np.uint32)
    out = {
        "sets": [tuple(int(x) for x in set_ids[int(set_ptr[i]):int(set_ptr[i + 1])]) for i in range(v.n_sets)],
        "posteriors": arr(v.posteriors, v.n_posteriors, np.float64),
        "abundances": arr(v.abundances, v.n_abundances, np.float64),
        "noise_count": float(v.noise_count),
        "total_count": float(v.total_count),
        "gibbs": [],
    }
    if v.n_gibbs:
        pp = arr(v.gibbs_path_ptr, v.n_gibbs + 1, np.uint64)
        npn = arr(v.gibbs_noise_ptr, v.n_gibbs + 1, np.uint64)
        ap = arr(v.gibbs_abund_ptr, v.n_gibbs + 1, np.uint64)
        pids = arr(v.gibbs_path_ids, int(pp[-1]), np.uint32)
        noise = arr(v.gibbs_noise, int(npn[-1]), np.float64)
        ab = arr(v.gibbs_abund, int(ap[-1]), np.float64)
        for g in range(v.n_gibbs):
            ids = pids[int(pp[g]):int(pp[g + 1])]
            ns = noise[int(npn[g]):int(npn[g + 1])]
            out["gibbs"].append({"path_ids": ids, "noise_samples": ns,
                                 "abundance_samples": ab[int(ap[g]):int(ap[g + 1])].reshape(len(ns), len(ids))})
    if free:
        lib_.harness_result_free(handle)
    return out


def harness_estimate(lib_, config: HarnessConfig, arena, c: int = 0, seed: int = 1) -> dict:
    """``PathEstimator::estimate`` of the model in ``config`` on cluster c, through a harness library (reference or ours)."""
    est = lib_.harness_estimator_create(C.byref(config))
    try:
        data = HarnessClusterData(arena, c)
        handle = lib_.harness_estimate(est, C.byref(data.struct), seed)
        return result_to_dict(lib_, handle)
    finally:
        lib_.harness_estimator_destroy(est)


def harness_estimate_batch(lib_, config: HarnessConfig, arena, cluster_ids=None, rng_seed: int = 1, num_threads: int = 0):
    """The inference loop (rpvg:src/main.cpp:811-830,976-977) over the given clusters. Returns (seconds, [result dicts])."""
    ids = list(range(arena.n_clusters)) if cluster_ids is None else list(cluster_ids)
    datas = [HarnessClusterData(arena, c) for c in ids]
    structs = (HarnessCluster * len(ids))(*[d.struct for d in datas])
    handles = (C.c_void_p * len(ids))()
    est = lib_.harness_estimator_create(C.byref(config))
    try:
        seconds = lib_.harness_estimate_batch(est, len(ids), structs, rng_seed, num_threads, handles)
        return seconds, [result_to_dict(lib_, handles[i]) for i in range(len(ids))]
    finally:
        lib_.harness_estimator_destroy(est)


def harness_estimate_batched(lib_, config: HarnessConfig, arena, cluster_ids=None, rng_seed: int = 1, num_devices: int = 1):
    """rpvg_b200 harness only: the same clusters, order and generators through the classes' estimateBatch (one call), or,
    with num_devices != 1, through estimateBatchOnDevices (the batch dealt over that many GPUs)."""
    ids = list(range(arena.n_clusters)) if cluster_ids is None else list(cluster_ids)
    datas = [HarnessClusterData(arena, c) for c in ids]
    structs = (HarnessCluster * len(ids))(*[d.struct for d in datas])
    handles = (C.c_void_p * len(ids))()
    lib_.harness_estimate_batched_devices.restype = C.c_double
    lib_.harness_estimate_batched_devices.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(HarnessCluster), C.c_uint32, C.c_int32, C.POINTER(C.c_void_p)]
    est = lib_.harness_estimator_create(C.byref(config))
    try:
        seconds = lib_.harness_estimate_batched_devices(est, len(ids), structs, rng_seed, num_devices, handles)
        return seconds, [result_to_dict(lib_, handles[i]) for i in range(len(ids))]
    finally:
        lib_.harness_estimator_destroy(est)




def harness_source_groups(lib, id_lists):
    """findPathSourceGroups of the haplotype-transcripts class on paths carrying the given haplotype ids (one list per path,
    inserted in list order). Returns (groups: list of path-index lists in output order, counts)."""
    n = len(id_lists)
    id_ptr = np.zeros(n + 1, dtype=np.uint64)
    id_ptr[1:] = np.cumsum([len(x) for x in id_lists])
    ids = np.asarray([i for x in id_lists for i in x], dtype=np.uint32)
    total = max(1, int(id_ptr[-1]))
    group_ptr = np.zeros(total + 1, dtype=np.uint64)
    group_paths = np.zeros(total, dtype=np.uint32)
    group_counts = np.zeros(total, dtype=np.uint32)
    g = lib.harness_source_groups(n, id_ptr.ctypes.data, ids.ctypes.data if ids.size else None, group_ptr.ctypes.data, group_paths.ctypes.data, group_counts.ctypes.data)
    groups = [[int(p) for p in group_paths[int(group_ptr[k]):int(group_ptr[k + 1])]] for k in range(g)]
    return groups, [int(c) for c in group_counts[:g]]
