"""The pair log-likelihood sums of rpvg_b200_grouped_bounded_pairs (rpvg:src/path_estimator.cpp:424-440: for every pair of
paths, sum_r count_r * log(noise_r + col_a / 2 + col_b / 2)) against a long-double evaluation of the same sums, on read
classes that take every route of the kernel: counts 1, small counts, counts above 31, noise far below the probabilities,
noise zero, and more read classes than one work item takes."""

import ctypes as C

import numpy as np
import pytest

from rpvg_b200 import capi
from rpvg_b200.arena import ClusterArena

pytestmark = pytest.mark.gpu


def make_cluster(rng, n_paths, n_rows, counts, noises, density=0.8):
    row_ptr, col, prob = [0], [], []
    for _ in range(n_rows):
        present = np.flatnonzero(rng.random(n_paths) < density)
        if present.size == 0:
            present = np.array([int(rng.integers(n_paths))])
        # probabilities over twelve decimal orders
        values = 10.0 ** rng.uniform(-12, 0, size=present.size)
        col.extend(int(j) for j in present)
        prob.extend(float(v) for v in values)
        row_ptr.append(len(col))
    count = rng.choice(np.asarray(counts, dtype=np.uint32), size=n_rows)
    noise = rng.choice(np.asarray(noises, dtype=np.float64), size=n_rows)
    return row_ptr, col, prob, count, noise


def arena_of(clusters, n_paths):
    read_offset, path_offset, row_ptr, col, prob, count, noise = [0], [0], [0], [], [], [], []
    for (rp, cl, pr, cn, nz), npaths in zip(clusters, n_paths):
        base = row_ptr[-1]
        row_ptr.extend(base + x for x in rp[1:])
        col.extend(cl)
        prob.extend(pr)
        count.extend(cn)
        noise.extend(nz)
        read_offset.append(read_offset[-1] + len(rp) - 1)
        path_offset.append(path_offset[-1] + npaths)
    return ClusterArena(np.asarray(read_offset, np.uint64), np.asarray(path_offset, np.uint64), np.asarray(row_ptr, np.uint64), np.asarray(col, np.uint32),
                        np.asarray(prob, np.float64), np.asarray(noise, np.float64), np.asarray(count, np.uint32))


def pair_sums_on_device(arena):
    lib = capi.lib()
    lib.rpvg_b200_grouped_raw.argtypes = [C.c_void_p, C.POINTER(capi.Clusters), C.POINTER(C.c_void_p)]
    lib.rpvg_b200_grouped_bounded_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double] + [C.POINTER(C.c_void_p)] * 3
    lib.rpvg_b200_grouped_free.argtypes = [C.c_void_p, C.c_void_p]

    ctx = capi.Context(0)
    cl = capi.clusters_struct(arena)
    handle = C.c_void_p()
    ctx._check(lib.rpvg_b200_grouped_raw(ctx._h, C.byref(cl), C.byref(handle)))

    group_offset = np.ascontiguousarray(arena.path_offset, dtype=np.uint64)
    order = np.concatenate([np.arange(int(arena.path_offset[c + 1] - arena.path_offset[c]), dtype=np.uint32) for c in range(arena.n_clusters)])
    log_freq = np.zeros(arena.n_paths, dtype=np.float64)
    outs = [C.c_void_p(), C.c_void_p(), C.c_void_p()]
    # nothing is pruned: every pair comes back as a candidate
    ctx._check(lib.rpvg_b200_grouped_bounded_pairs(ctx._h, handle, group_offset.ctypes.data, order.ctypes.data, log_freq.ctypes.data, float("-inf"),
                                                   C.byref(outs[0]), C.byref(outs[1]), C.byref(outs[2])))

    result = []
    offsets = np.ctypeslib.as_array(C.cast(outs[0], C.POINTER(C.c_uint64)), shape=(arena.n_clusters + 1,)).copy()
    total = int(offsets[-1])
    pairs = np.ctypeslib.as_array(C.cast(outs[1], C.POINTER(C.c_uint32)), shape=(total,)).copy()
    sums = np.ctypeslib.as_array(C.cast(outs[2], C.POINTER(C.c_double)), shape=(total,)).copy()
    for c in range(arena.n_clusters):
        result.append((pairs[int(offsets[c]):int(offsets[c + 1])], sums[int(offsets[c]):int(offsets[c + 1])]))

    lib.rpvg_b200_grouped_free(ctx._h, handle)
    ctx.close()
    return result


def pair_sums_long_double(arena, c):
    r0, r1 = int(arena.read_offset[c]), int(arena.read_offset[c + 1])
    n_paths = int(arena.path_offset[c + 1] - arena.path_offset[c])
    dense = np.zeros((r1 - r0, n_paths), dtype=np.float64)
    for r in range(r0, r1):
        e0, e1 = int(arena.row_ptr[r]), int(arena.row_ptr[r + 1])
        dense[r - r0, arena.col[e0:e1]] = arena.prob[e0:e1]
    noise = arena.noise[r0:r1]
    count = arena.count[r0:r1].astype(np.longdouble)
    out = []
    for i in range(n_paths):
        for j in range(i, n_paths):
            # the double-precision terms of the reference, their logs and the sum in extended precision
            terms = (noise + dense[:, i] / 2.0) + dense[:, j] / 2.0
            with np.errstate(divide="ignore"):
                out.append(float(np.sum(count * np.log(terms.astype(np.longdouble)))))
    return np.asarray(out)


@pytest.mark.parametrize("seed", [1, 2])
def test_pair_sums_every_route(seed):
    rng = np.random.default_rng(seed)
    clusters, n_paths = [], []
    # all counts 1, ordinary noise; small counts; counts beyond the powers; noise far below / zero; a long cluster
    specs = [(40, 300, [1], [1e-4]),
             (37, 500, [1, 2, 3, 4, 5, 6, 7, 9, 16, 31], [1e-4, 1e-3, 0.25]),
             (70, 200, [32, 33, 100, 1000, 4119, 250000], [1e-4]),
             (20, 400, [1, 2, 50], [1e-30, 1e-200, 0.0, 1e-4]),
             (5, 40000, [1, 1, 1, 2, 3, 40], [1e-4, 1e-6])]
    for paths, rows, counts, noises in specs:
        clusters.append(make_cluster(rng, paths, rows, counts, noises))
        n_paths.append(paths)
    arena = arena_of(clusters, n_paths)

    device = pair_sums_on_device(arena)

    for c in range(arena.n_clusters):
        pairs, sums = device[c]
        expect = pair_sums_long_double(arena, c)
        assert np.array_equal(pairs, np.arange(expect.size, dtype=np.uint32)), "every pair is a candidate, in enumeration order"
        finite = np.isfinite(expect)
        assert np.array_equal(np.isfinite(sums), finite)
        assert np.array_equal(sums[~finite], expect[~finite])   # -inf where a term is zero
        # one rounding of a sum of this size is 1.1e-16 relative; the products lose a little more on long clusters
        rel = np.max(np.abs(sums[finite] - expect[finite]) / np.abs(expect[finite]), initial=0)
        assert rel < 2e-13, (c, rel)
