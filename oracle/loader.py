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

# the harness ABI is shared with this repository's own build of the harness: its binding lives with the harness
from tests.harness.binding import (HarnessCluster, HarnessConfig, HarnessResult, HarnessClusterData, MODEL_TRANSCRIPTS, MODEL_STRAINS,  # noqa: F401
                                   MODEL_HAPLOTYPE_TRANSCRIPTS, MODEL_PATH_POSTERIOR, MODEL_HAPLOTYPES, make_config, load_harness, result_to_dict,
                                   harness_estimate, harness_estimate_batch, harness_estimate_batched)

_HERE = os.path.dirname(os.path.abspath(__file__))

_port = None
_reference = None


class OracleClusters(C.Structure):
    _fields_ = [("n_clusters", C.c_uint32)] + [(n, C.c_void_p) for n in
                                               ("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count")]


class HarnessAlignments(C.Structure):
    _fields_ = [("num_paths", C.c_uint32), ("effective_length", C.c_void_p), ("path_group", C.c_void_p), ("source_count", C.c_void_p),
                ("num_groups", C.c_uint32), ("num_reads", C.c_uint32), ("min_mapq", C.c_void_p), ("align_ptr", C.c_void_p), ("score_sum", C.c_void_p),
                ("align_length", C.c_void_p), ("frag_length", C.c_void_p), ("id_ptr", C.c_void_p), ("ids", C.c_void_p)]


class HarnessEstimates(C.Structure):
    _fields_ = [("cluster_id", C.c_uint32), ("num_paths", C.c_uint32), ("length", C.c_void_p), ("effective_length", C.c_void_p),
                ("n_sets", C.c_uint64), ("set_ptr", C.c_void_p), ("set_ids", C.c_void_p), ("posteriors", C.c_void_p),
                ("n_abundances", C.c_uint64), ("abundances", C.c_void_p), ("noise_count", C.c_double), ("total_count", C.c_double)]


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
