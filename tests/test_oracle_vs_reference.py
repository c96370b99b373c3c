"""Pins the CPU port (oracle/rpvg_oracle.cpp) against the reference's own compiled code (oracle/_ref) routine by routine.

CPU only. Skipped where oracle/_ref has not been built (it needs /root/reference at build time; the built library ships).
"""

import ctypes as C

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import synth


def dense_cluster(rng, rows, n_paths, density=0.3, normalised=True):
    """Random dense col-major matrix as addNoiseAndNormalize leaves it (rows sum to 1 with the noise column)."""
    mat = rng.random((rows, n_paths)) * (rng.random((rows, n_paths)) < density)
    noise = np.maximum(1e-4, 10.0 ** (-rng.choice([60, 30, 10, 1], size=rows) / 10.0))
    counts = rng.geometric(0.3, size=rows).astype(np.float64)
    if normalised:
        s = mat.sum(axis=1, keepdims=True)
        with np.errstate(invalid="ignore", divide="ignore"):
            mat = np.where(s > 0, mat / s * (1 - noise[:, None]), 0.0)
    return np.asfortranarray(mat), noise, counts


def ref_posteriors(ref, kind, k, min_rel, seed, mat, noise, counts, path_counts):
    rows, n = mat.shape
    h = ref.ref_group_posteriors(kind, k, min_rel, seed, rows, n, mat.ctypes.data, noise.ctypes.data, counts.ctypes.data, path_counts.ctypes.data)
    return loader.result_to_dict(ref, h)


def port_posteriors(port, kind, k, min_rel, seed, mat, noise, counts, path_counts, cap=200000):
    rows, n = mat.shape
    sets = np.zeros(cap * k, dtype=np.uint32)
    post = np.zeros(cap, dtype=np.float64)
    got = port.oracle_group_posteriors(kind, k, min_rel, seed, rows, n, mat.ctypes.data, noise.ctypes.data, counts.ctypes.data, path_counts.ctypes.data,
                                       cap, sets.ctypes.data, post.ctypes.data)
    assert got >= 0
    return [tuple(int(x) for x in sets[i * k:(i + 1) * k]) for i in range(got)], post[:got]


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_em_port_matches_reference_estimator(port, reference, seed):
    arena = synth.transcripts_arena(seed=seed, n_paths=1200, n_reads=6000, n_max=150)
    ab, noise_count, total, its, _ = loader.port_em_abundances(arena)
    _, results = loader.harness_estimate_batch(reference, loader.make_config(loader.MODEL_TRANSCRIPTS), arena)
    for c, res in enumerate(results):
        p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
        mine, ref = ab[p0:p1], res["abundances"]
        assert np.array_equal(mine == 0, ref == 0)
        nz = ref != 0
        if nz.any():
            assert np.max(np.abs(mine[nz] - ref[nz]) / ref[nz]) < 1e-11
        assert abs(noise_count[c] - res["noise_count"]) <= 1e-10 * max(1.0, total[c])
        assert total[c] == res["total_count"]


def test_em_dense_port_matches_ref_em(port, reference):
    rng = np.random.default_rng(5)
    for rows, n in [(40, 3), (300, 12), (1500, 60)]:
        mat, noise, counts = dense_cluster(rng, rows, n)
        full = np.asfortranarray(np.hstack([mat, noise[:, None]]))
        ab_r = np.zeros(n); nc_r = C.c_double(); tc_r = C.c_double()
        reference.ref_em(10000, 0.001, rows, n, full.ctypes.data, counts.ctypes.data, ab_r.ctypes.data, C.byref(nc_r), C.byref(tc_r))
        ab_p = np.zeros(n); nc_p = C.c_double(); tc_p = C.c_double()
        its = port.oracle_em_dense(10000, 0.001, rows, n, full.ctypes.data, counts.ctypes.data, ab_p.ctypes.data, C.byref(nc_p), C.byref(tc_p))
        assert its >= 10
        assert np.array_equal(ab_r == 0, ab_p == 0)
        assert np.allclose(ab_r, ab_p, rtol=1e-11, atol=0)
        assert abs(nc_r.value - nc_p.value) <= 1e-10 * tc_r.value and tc_r.value == tc_p.value


def test_collapse_port_is_bit_identical_to_reference(port, reference):
    """readCollapseProbabilityMatrix: same classes, counts, order and representative values, bit for bit — also with
    near-duplicate rows (1-2 ulp apart) and rows closer than prob_precision."""
    rng = np.random.default_rng(11)
    for rows, cols in [(50, 4), (400, 7), (2000, 5)]:
        base = rng.random((rows // 4, cols))
        base /= base.sum(axis=1, keepdims=True)
        pick = rng.integers(0, len(base), size=rows)
        mat = base[pick].copy()
        jitter = rng.integers(0, 4, size=rows)
        mat[jitter == 1] *= (1 + 2.2e-16)          # comparator-equal, bitwise different
        mat[jitter == 2, 0] += 3e-9                # inside prob_precision, outside double_precision
        mat[jitter == 3, 1] += 5e-8                # outside prob_precision
        mat = np.asfortranarray(mat)
        counts = rng.geometric(0.3, size=rows).astype(np.float64)

        out_r = np.zeros((rows, cols), order="F"); cnt_r = np.zeros(rows)
        n_r = reference.ref_collapse(1e-8, rows, cols, mat.ctypes.data, counts.ctypes.data, out_r.ctypes.data, cnt_r.ctypes.data)
        out_p = np.zeros(rows * cols); cnt_p = np.zeros(rows); cls = np.zeros(rows, dtype=np.uint32)
        n_p = port.oracle_collapse_dense(1e-8, rows, cols, mat.ctypes.data, counts.ctypes.data, out_p.ctypes.data, cnt_p.ctypes.data, cls.ctypes.data)

        assert n_r == n_p and n_r < rows
        ref_mat = out_r.ravel(order="F")[:n_r * cols].reshape((n_r, cols), order="F")
        port_mat = out_p[:n_p * cols].reshape((n_p, cols), order="F")
        assert np.array_equal(ref_mat, port_mat)
        assert np.array_equal(cnt_r[:n_r], cnt_p[:n_p])
        assert np.array_equal(np.bincount(cls, weights=counts, minlength=n_p), cnt_p[:n_p])


@pytest.mark.parametrize("kind,k", [(0, 1), (0, 2), (0, 3), (1, 2)])
def test_group_posteriors_port_matches_reference(port, reference, kind, k):
    rng = np.random.default_rng(21 + kind * 10 + k)
    for rows, n in [(60, 4), (500, 9)]:
        mat, noise, counts = dense_cluster(rng, rows, n, density=0.5)
        path_counts = rng.integers(1, 6, size=n).astype(np.uint32)
        min_rel = 1e-8 if kind == 1 else 0.0
        ref = ref_posteriors(reference, kind, k, min_rel, 0, mat, noise, counts, path_counts)
        sets, post = port_posteriors(port, kind, k, min_rel, 0, mat, noise, counts, path_counts)
        assert ref["sets"] == sets
        assert np.allclose(ref["posteriors"], post, rtol=0, atol=1e-12)
        assert abs(post.sum() - 1) < 1e-9


def test_group_posteriors_gibbs_port_replays_the_reference_chain(port, reference):
    """Same mt19937 seed, same libstdc++ distributions, same consumption order -> identical visit counts."""
    rng = np.random.default_rng(33)
    mat, noise, counts = dense_cluster(rng, 200, 6, density=0.6)
    path_counts = rng.integers(1, 4, size=6).astype(np.uint32)
    for k, seed in [(2, 7), (3, 8), (4, 9)]:
        ref = ref_posteriors(reference, 2, k, 0.0, seed, mat, noise, counts, path_counts)
        sets, post = port_posteriors(port, 2, k, 0.0, seed, mat, noise, counts, path_counts)
        assert ref["sets"] == sets
        assert np.allclose(ref["posteriors"], post, rtol=0, atol=1e-12)


def test_path_cover_known_answer_and_reference(port, reference):
    """The reference's only known-answer test on the path (rpvg:src/tests/path_abundance_estimator_test.cpp:8-28)."""
    cover = np.asfortranarray(np.array([[1, 0, 0], [1, 1, 0], [0, 1, 1], [0, 1, 0]], dtype=np.uint8))
    counts = np.array([1.0, 1.0, 1.0, 1.0])
    for weights, expect in [(np.array([1.0, 1.0, 1.0]), [0, 1]), (np.array([1.0, 1.0, 0.01]), [0, 1, 2])]:
        for fn in (reference.ref_path_cover, port.oracle_path_cover):
            out = np.zeros(3, dtype=np.uint32)
            n = fn(4, 3, cover.ctypes.data, counts.ctypes.data, weights.ctypes.data, out.ctypes.data)
            assert list(out[:n]) == expect
    rng = np.random.default_rng(3)
    for rows, n in [(30, 6), (200, 25)]:
        cov = np.asfortranarray((rng.random((rows, n)) < 0.25).astype(np.uint8))
        cov[np.arange(rows), rng.integers(0, n, size=rows)] = 1
        cnt = rng.geometric(0.3, size=rows).astype(np.float64)
        w = rng.random(n) * 50 + 1
        a = np.zeros(n, dtype=np.uint32); b = np.zeros(n, dtype=np.uint32)
        na = reference.ref_path_cover(rows, n, cov.ctypes.data, cnt.ctypes.data, w.ctypes.data, a.ctypes.data)
        nb = port.oracle_path_cover(rows, n, cov.ctypes.data, cnt.ctypes.data, w.ctypes.data, b.ctypes.data)
        assert na == nb and np.array_equal(a[:na], b[:nb])


def test_gibbs_read_counts_port_replays_the_reference_chain(port, reference):
    rng = np.random.default_rng(41)
    rows, n, samples, thin = 120, 5, 6, 5
    mat, noise, counts = dense_cluster(rng, rows, n, density=0.6)
    full = np.asfortranarray(np.hstack([mat, noise[:, None]]))
    ab = np.zeros(n); nc = C.c_double(); ns_r = np.zeros(samples); as_r = np.zeros(samples * n)
    reference.ref_em_gibbs(10000, 0.001, samples, thin, 5, rows, n, full.ctypes.data, counts.ctypes.data, ab.ctypes.data, C.byref(nc),
                           ns_r.ctypes.data, as_r.ctypes.data)
    ns_p = np.zeros(samples); as_p = np.zeros(samples * n)
    port.oracle_gibbs_read_counts(samples, thin, 5, rows, n, full.ctypes.data, counts.ctypes.data, ab.ctypes.data, nc.value, counts.sum(),
                                  ns_p.ctypes.data, as_p.ctypes.data)
    assert np.allclose(ns_r, ns_p, rtol=1e-12) and np.allclose(as_r, as_p, rtol=1e-12, atol=0)
    assert np.allclose(as_p.reshape(samples, n).sum(axis=1) + ns_p, counts.sum(), rtol=1e-9)
