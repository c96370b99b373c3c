"""TEST INFRASTRUCTURE — ctypes loaders for the two CPU oracles. NOT product code.

* ``port()``      -> oracle/liboracle.so: our CPU restatement (oracle/rpvg_oracle.cpp).
* ``reference()`` -> oracle/_ref/libharness_ref*.so: rpvg's own unmodified translation units behind
  tests/harness/estimator_harness.h (built by oracle/Makefile in the container where /root/reference exists; the
  built file travels to the GPU box, the sources do not).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline / ``--impl reference`` legs import this module.
"""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

_port = None
_reference = None


class OracleClusters(C.Structure):
    _fields_ = [("n_clusters", C.c_uint32)] + [(n, C.c_void_p) for n in
                                               ("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count")]


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


class HarnessAlignments(C.Structure):
    _fields_ = [("num_paths", C.c_uint32), ("effective_length", C.c_void_p), ("path_group", C.c_void_p), ("source_count", C.c_void_p),
                ("num_groups", C.c_uint32), ("num_reads", C.c_uint32), ("min_mapq", C.c_void_p), ("align_ptr", C.c_void_p), ("score_sum", C.c_void_p),
                ("align_length", C.c_void_p), ("frag_length", C.c_void_p), ("id_ptr", C.c_void_p), ("ids", C.c_void_p)]


class HarnessEstimates(C.Structure):
    _fields_ = [("cluster_id", C.c_uint32), ("num_paths", C.c_uint32), ("length", C.c_void_p), ("effective_length", C.c_void_p),
                ("n_sets", C.c_uint64), ("set_ptr", C.c_void_p), ("set_ids", C.c_void_p), ("posteriors", C.c_void_p),
                ("n_abundances", C.c_uint64), ("abundances", C.c_void_p), ("noise_count", C.c_double), ("total_count", C.c_double)]


MODEL_TRANSCRIPTS, MODEL_STRAINS, MODEL_HAPLOTYPE_TRANSCRIPTS, MODEL_PATH_POSTERIOR, MODEL_HAPLOTYPES = range(5)


def make_config(model, group_size=2, min_hap_prob=0.001, infer_collapsed=True, use_group_post_gibbs=False, max_em_its=10000,
                max_rel_em_conv=0.001, num_gibbs_samples=0, gibbs_thin_its=25, prob_precision=1e-8) -> HarnessConfig:
    """Defaults are rpvg's command-line defaults (rpvg:src/main.cpp:402-418)."""
    return HarnessConfig(model, group_size, min_hap_prob, int(infer_collapsed), int(use_group_post_gibbs), max_em_its,
                         max_rel_em_conv, num_gibbs_samples, gibbs_thin_its, prob_precision)


def port_path() -> str:
    return os.path.join(_HERE, "liboracle.so")


def port():
    global _port
    if _port is None:
        l = C.CDLL(port_path())
        l.oracle_max_threads.restype = C.c_int32
        l.oracle_em_abundances.restype = C.c_double
        l.oracle_em_abundances.argtypes = [C.POINTER(OracleClusters), C.c_uint32, C.c_double, C.c_uint32, C.c_double, C.c_int32] + [C.c_void_p] * 4
        l.oracle_normalize.argtypes = [C.c_uint32] + [C.c_void_p] * 4
        l.oracle_em_dense.restype = C.c_uint32
        l.oracle_em_dense.argtypes = [C.c_uint32, C.c_double, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5
        l.oracle_collapse_dense.restype = C.c_uint32
        l.oracle_collapse_dense.argtypes = [C.c_double, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5
        l.oracle_group_posteriors.restype = C.c_int64
        l.oracle_group_posteriors.argtypes = [C.c_int32, C.c_uint32, C.c_double, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 4 + [C.c_uint64, C.c_void_p, C.c_void_p]
        l.oracle_path_cover.restype = C.c_uint32
        l.oracle_path_cover.argtypes = [C.c_uint32, C.c_uint32] + [C.c_void_p] * 4
        l.oracle_gibbs_read_counts.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        _port = l
    return _port


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def reference_path() -> str | None:
    """The -march=native build if this CPU has every flag of the CPU it was built on, else the x86-64-v3 build."""
    ref_dir = os.path.join(_HERE, "_ref")
    native = os.path.join(ref_dir, "libharness_ref_native.so")
    flags_file = os.path.join(ref_dir, "native_cpu_flags.txt")
    if os.path.exists(native) and os.path.exists(flags_file) and not os.environ.get("RPVG_ORACLE_NO_NATIVE"):
        with open(flags_file) as f:
            built = set(f.read().split(":", 1)[1].split())
        if built and built <= _cpu_flags():
            return native
    generic = os.path.join(ref_dir, "libharness_ref.so")
    return generic if os.path.exists(generic) else None


def reference_available() -> bool:
    return reference_path() is not None


def _bind_harness(l):
    l.harness_impl_name.restype = C.c_char_p
    l.harness_max_threads.restype = C.c_int32
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


def reference():
    global _reference
    if _reference is None:
        path = reference_path()
        if path is None:
            raise RuntimeError("oracle/_ref/libharness_ref.so is missing: run `make -C oracle ref` where /root/reference exists")
        l = load_harness(path)
        l.ref_build_matrix.argtypes = [C.c_double, C.POINTER(HarnessCluster), C.c_int32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int32] + [C.c_void_p] * 3
        l.ref_merge_read_records.restype = C.c_uint32
        l.ref_merge_read_records.argtypes = [C.c_double, C.POINTER(HarnessCluster)] + [C.c_void_p] * 6
        l.ref_collapse.restype = C.c_uint32
        l.ref_collapse.argtypes = [C.c_double, C.c_uint32, C.c_uint32] + [C.c_void_p] * 4
        l.ref_em.argtypes = [C.c_uint32, C.c_double, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5
        l.ref_em_gibbs.argtypes = [C.c_uint32, C.c_double, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 6
        l.ref_group_posteriors.restype = C.c_void_p
        l.ref_group_posteriors.argtypes = [C.c_int32, C.c_uint32, C.c_double, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 4
        l.ref_path_cover.restype = C.c_uint32
        l.ref_path_cover.argtypes = [C.c_uint32, C.c_uint32] + [C.c_void_p] * 4
        l.ref_frag_log_prob.restype = C.c_double
        l.ref_frag_log_prob.argtypes = [C.c_double, C.c_double, C.c_uint32, C.c_uint32]
        l.ref_add_path_probs.restype = C.c_uint64
        l.ref_add_path_probs.argtypes = [C.POINTER(HarnessAlignments), C.c_double, C.c_double, C.c_uint32, C.c_int32, C.c_double, C.c_double, C.c_int32] + [C.c_void_p] * 5
        l.ref_write_probs.restype = C.c_int32
        l.ref_write_probs.argtypes = [C.c_char_p, C.c_double, C.c_uint32, C.POINTER(HarnessCluster), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        l.ref_write_estimates.restype = C.c_int32
        l.ref_write_estimates.argtypes = [C.c_int32, C.c_char_p, C.c_uint32, C.c_double, C.c_double, C.c_uint32, C.POINTER(HarnessEstimates), C.c_uint32]
        _reference = l
    return _reference


# ---- marshalling ------------------------------------------------------------------------------------------------------

def oracle_clusters(arena) -> OracleClusters:
    return OracleClusters(arena.n_clusters, arena.read_offset.ctypes.data, arena.path_offset.ctypes.data, arena.row_ptr.ctypes.data,
                          arena.col.ctypes.data, arena.prob.ctypes.data, arena.noise.ctypes.data, arena.count.ctypes.data)


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


def harness_estimate_batched(lib_, config: HarnessConfig, arena, cluster_ids=None, rng_seed: int = 1):
    """rpvg_b200 harness only: the same clusters, order and generators through the classes' estimateBatch (one call)."""
    ids = list(range(arena.n_clusters)) if cluster_ids is None else list(cluster_ids)
    datas = [HarnessClusterData(arena, c) for c in ids]
    structs = (HarnessCluster * len(ids))(*[d.struct for d in datas])
    handles = (C.c_void_p * len(ids))()
    lib_.harness_estimate_batched.restype = C.c_double
    lib_.harness_estimate_batched.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(HarnessCluster), C.c_uint32, C.POINTER(C.c_void_p)]
    est = lib_.harness_estimator_create(C.byref(config))
    try:
        seconds = lib_.harness_estimate_batched(est, len(ids), structs, rng_seed, handles)
        return seconds, [result_to_dict(lib_, handles[i]) for i in range(len(ids))]
    finally:
        lib_.harness_estimator_destroy(est)


# ---- convenience wrappers over the port ---------------------------------------------------------------------------------

def port_em_abundances(arena, max_em_its=10000, max_rel_em_conv=0.001, min_em_conv_its=10, min_em_abundance=1e-8, num_threads=0):
    """Returns (abundances, noise_count, total_count, iterations, seconds)."""
    abundances = np.empty(arena.n_paths, dtype=np.float64)
    noise_count = np.empty(arena.n_clusters, dtype=np.float64)
    total_count = np.empty(arena.n_clusters, dtype=np.float64)
    iterations = np.empty(arena.n_clusters, dtype=np.uint32)
    cl = oracle_clusters(arena)
    seconds = port().oracle_em_abundances(C.byref(cl), max_em_its, max_rel_em_conv, min_em_conv_its, min_em_abundance, num_threads,
                                          abundances.ctypes.data, noise_count.ctypes.data, total_count.ctypes.data, iterations.ctypes.data)
    return abundances, noise_count, total_count, iterations, seconds


def ref_merge_read_records(ref, num_paths, count, noise, entry_ptr, entry_prob, entry_path_ptr, entry_path_ids, prob_precision=1e-8):
    """The reference's own sort + quickMergeIdentical scan (rpvg:src/main.cpp:953-973) over one cluster's records.
    -> (noise, count, entry_ptr, entry_prob, entry_path_ptr, entry_path_ids) of the kept records."""
    count = np.ascontiguousarray(count, dtype=np.uint32)
    noise = np.ascontiguousarray(noise, dtype=np.float64)
    entry_ptr = np.ascontiguousarray(entry_ptr, dtype=np.uint64)
    entry_prob = np.ascontiguousarray(entry_prob, dtype=np.float64)
    entry_path_ptr = np.ascontiguousarray(entry_path_ptr, dtype=np.uint64)
    entry_path_ids = np.ascontiguousarray(entry_path_ids, dtype=np.uint32)
    R = len(count)
    cluster = HarnessCluster(num_paths, R, count.ctypes.data, noise.ctypes.data, entry_ptr.ctypes.data, entry_prob.ctypes.data, entry_path_ptr.ctypes.data,
                             entry_path_ids.ctypes.data, None, None, None, None)
    o_noise = np.zeros(R, dtype=np.float64)
    o_count = np.zeros(R, dtype=np.uint32)
    o_entry_ptr = np.zeros(R + 1, dtype=np.uint64)
    o_entry_prob = np.zeros(len(entry_prob) + 1, dtype=np.float64)
    o_path_ptr = np.zeros(len(entry_prob) + 1, dtype=np.uint64)
    o_ids = np.zeros(len(entry_path_ids) + 1, dtype=np.uint32)
    n = ref.ref_merge_read_records(prob_precision, C.byref(cluster), o_noise.ctypes.data, o_count.ctypes.data, o_entry_ptr.ctypes.data, o_entry_prob.ctypes.data,
                                   o_path_ptr.ctypes.data, o_ids.ctypes.data)
    ne = int(o_entry_ptr[n])
    return o_noise[:n], o_count[:n], o_entry_ptr[:n + 1], o_entry_prob[:ne], o_path_ptr[:ne + 1], o_ids[:int(o_path_ptr[ne])]


def ref_add_path_probs(ref, effective_length, min_mapq, align_ptr, score_sum, align_length, frag_length, id_ptr, ids, frag_mean=10.0, frag_sd=2.0,
                       sd_max_multi=10, is_single_end=False, min_noise_prob=0.0, prob_precision=1e-8, path_group=None, source_count=None, num_groups=0):
    """The reference's own ReadPathProbabilities::addPathProbs (rpvg:src/read_path_probabilities.cpp:74-221) on the reads of one
    cluster. -> (noise, group_ptr, group_prob, group_id_ptr, group_ids)."""
    def arr(a, t):
        return None if a is None else np.ascontiguousarray(a, dtype=t)

    def p(a):
        return None if a is None else a.ctypes.data

    effective_length = arr(effective_length, np.float64)
    min_mapq, score_sum = arr(min_mapq, np.uint8), arr(score_sum, np.int32)
    align_ptr, id_ptr = arr(align_ptr, np.uint64), arr(id_ptr, np.uint64)
    align_length, frag_length, ids = arr(align_length, np.uint32), arr(frag_length, np.uint32), arr(ids, np.uint32)
    path_group, source_count = arr(path_group, np.uint32), arr(source_count, np.uint32)
    R, T = len(min_mapq), len(ids)
    al = HarnessAlignments(len(effective_length), p(effective_length), p(path_group), p(source_count), num_groups, R, p(min_mapq), p(align_ptr), p(score_sum),
                           p(align_length), p(frag_length), p(id_ptr), p(ids))
    noise = np.zeros(R, dtype=np.float64)
    group_ptr = np.zeros(R + 1, dtype=np.uint64)
    group_prob = np.zeros(T + 1, dtype=np.float64)
    group_id_ptr = np.zeros(T + 2, dtype=np.uint64)
    group_ids = np.zeros(T + 1, dtype=np.uint32)
    G = ref.ref_add_path_probs(C.byref(al), frag_mean, frag_sd, sd_max_multi, int(is_single_end), min_noise_prob, prob_precision, int(path_group is not None),
                               noise.ctypes.data, group_ptr.ctypes.data, group_prob.ctypes.data, group_id_ptr.ctypes.data, group_ids.ctypes.data)
    return noise, group_ptr, group_prob[:G], group_id_ptr[:G + 1], group_ids[:int(group_id_ptr[G])]


def ref_write_probs(ref, prefix, arena, lengths, effective_lengths, prob_precision=1e-8):
    """rpvg's own ProbabilityClusterWriter (rpvg:src/threaded_output_writer.cpp:40-95) on every cluster of the arena, in order
    -> ``<prefix>.txt.gz``. Path p of cluster c is named ``c<c>_p<p>``."""
    datas = [HarnessClusterData(arena, c) for c in range(arena.n_clusters)]
    structs = (HarnessCluster * arena.n_clusters)(*[d.struct for d in datas])
    lens, effs = [], []
    for c in range(arena.n_clusters):
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        lens.append(np.ascontiguousarray(lengths[p0:p1], dtype=np.uint32))
        effs.append(np.ascontiguousarray(effective_lengths[p0:p1], dtype=np.float64))
    len_ptrs = (C.c_void_p * arena.n_clusters)(*[a.ctypes.data for a in lens])
    eff_ptrs = (C.c_void_p * arena.n_clusters)(*[a.ctypes.data for a in effs])
    status = ref.ref_write_probs(os.fsencode(prefix), prob_precision, arena.n_clusters, structs, len_ptrs, eff_ptrs)
    assert status == 0
    return prefix + ".txt.gz"


def ref_write_estimates(ref, kind, prefix, clusters, ploidy=2, min_posterior=1e-8, total_transcript_count=1.0, unaligned_read_count=0):
    """rpvg's own estimate writers (rpvg:src/threaded_output_writer.cpp:237-548) -> ``<prefix>.txt``. kind 0: abundance table,
    1: haplotype abundance table, 2: joint haplotype abundance table, 3: joint haplotype table. ``clusters``: list of dicts
    with cluster_id, length, effective_length, sets (list of tuples), posteriors, abundances, noise_count, total_count."""
    keep, structs = [], []
    for cl in clusters:
        length = np.ascontiguousarray(cl["length"], dtype=np.uint32)
        eff = np.ascontiguousarray(cl["effective_length"], dtype=np.float64)
        set_ptr = np.zeros(len(cl["sets"]) + 1, dtype=np.uint64)
        set_ptr[1:] = np.cumsum([len(s) for s in cl["sets"]])
        set_ids = np.ascontiguousarray([x for s in cl["sets"] for x in s], dtype=np.uint32)
        post = np.ascontiguousarray(cl["posteriors"], dtype=np.float64)
        ab = np.ascontiguousarray(cl["abundances"], dtype=np.float64)
        keep.append((length, eff, set_ptr, set_ids, post, ab))
        structs.append(HarnessEstimates(cl["cluster_id"], len(length), length.ctypes.data, eff.ctypes.data, len(cl["sets"]), set_ptr.ctypes.data,
                                        set_ids.ctypes.data, post.ctypes.data, len(ab), ab.ctypes.data, float(cl["noise_count"]), float(cl["total_count"])))
    arr = (HarnessEstimates * len(structs))(*structs)
    status = ref.ref_write_estimates(kind, os.fsencode(prefix), ploidy, min_posterior, total_transcript_count, len(structs), arr, unaligned_read_count)
    assert status == 0
    return prefix + ".txt"
