"""rpvg_b200_em_cluster_noalloc — the entry point for callers that may not allocate device memory while running
(fixed memory budgets, real-time loops, hosts inside another allocator). The caller owns every buffer; the kernel's work
area is a fixed array inside the library. Here: two calls in a row on different clusters (the second finds the first
call's leftovers in the shared work area), each checked against the regular, allocating entry point through the same
C ABI and against the CPU oracle, with the tolerance of the other EM tests; plus the size limit and the on-device input
checks (a caller that understates nnz, or passes unsorted path indices, gets iterations = 0xffffffff and untouched
outputs).

Collected last on purpose (file name), and self-contained: it passes alone in a fresh process and after the other tests.
"""

import numpy as np
import pytest

from rpvg_b200 import capi, synth

pytestmark = pytest.mark.gpu

REL_TOL = 1e-6


def run_noalloc(torch, sub, nnz=None):
    dev = torch.device("cuda", 0)

    def up(a, dtype):
        return torch.from_numpy(a.view(dtype)).to(dev)

    row_ptr = up(sub.row_ptr, np.int64)
    col = up(sub.col, np.int32)
    prob = up(sub.prob, np.float64)
    noise = up(sub.noise, np.float64)
    count = up(sub.count, np.int32)
    abundances = torch.full((sub.n_paths,), -1.0, dtype=torch.float64, device=dev)
    scalars = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
    iterations = torch.zeros(1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    status, message = capi.em_cluster_noalloc(0, 0, sub.n_reads, sub.n_paths, sub.nnz_path if nnz is None else nnz, row_ptr.data_ptr(), col.data_ptr(), prob.data_ptr(), noise.data_ptr(),
                                              count.data_ptr(), capi.default_em_params(), abundances.data_ptr(), scalars.data_ptr(),
                                              scalars.data_ptr() + 8, iterations.data_ptr())
    assert status == 0, message
    torch.cuda.synchronize()
    return abundances.cpu().numpy(), float(scalars[0]), float(scalars[1]), int(iterations.cpu().numpy().view(np.uint32)[0])


def test_noalloc_entry_point_matches_regular_entry_point_twice_in_a_row():
    import torch
    from oracle import loader

    arena = synth.transcripts_arena(seed=77, n_paths=3000, n_reads=15000, n_max=400)
    weights = [arena.cluster_shape(c)[0] + arena.cluster_shape(c)[2] for c in range(arena.n_clusters)]
    big, small = int(np.argmax(weights)), int(np.argsort(weights)[len(weights) // 2])

    ctx = capi.Context(0)
    try:
        for c in (big, small):   # second call: a smaller cluster over the first call's leftovers
            sub = arena.select([c])
            ref_ab, ref_noise, ref_total, ref_its = ctx.em_abundances(sub)
            ab, noise_count, total_count, its = run_noalloc(torch, sub)

            assert total_count == ref_total[0]
            assert np.array_equal(ab == 0, ref_ab == 0)
            nz = ref_ab != 0
            assert np.max(np.abs(ab[nz] - ref_ab[nz]) / ref_ab[nz]) < REL_TOL
            assert abs(noise_count - ref_noise[0]) <= REL_TOL * max(1.0, ref_total[0])
            assert 0 < its <= 10000 and abs(its - int(ref_its[0])) <= 2   # atomics: summation order may differ from run to run

            # and against the CPU oracle (the regular entry point is itself checked against it in test_em_gpu.py)
            o_ab, o_noise, o_total, o_its, _ = loader.port_em_abundances(sub)
            assert total_count == o_total[0]
            assert np.array_equal(ab == 0, o_ab == 0)
            nz = o_ab != 0
            assert np.max(np.abs(ab[nz] - o_ab[nz]) / o_ab[nz]) < REL_TOL
            assert abs(noise_count - o_noise[0]) <= REL_TOL * max(1.0, o_total[0])
            assert abs(its - int(o_its[0])) <= 2
    finally:
        ctx.close()


def test_noalloc_entry_point_rejects_clusters_beyond_its_fixed_work_area():
    params = capi.default_em_params()
    status, message = capi.em_cluster_noalloc(0, 0, 10, 600000, 100, 8, 8, 8, 8, 8, params, 8, 8, 8, 8)   # never launched
    assert status == -4 and "fixed area holds" in message


def test_noalloc_entry_point_checks_its_inputs_on_the_device():
    import copy

    import torch

    arena = synth.transcripts_arena(seed=78, n_paths=400, n_reads=2000, n_max=60)
    weights = [arena.cluster_shape(c)[0] + arena.cluster_shape(c)[2] for c in range(arena.n_clusters)]
    sub = arena.select([int(np.argmax(weights))])
    assert sub.nnz_path > 8

    # nnz understated (the work area would be laid out past what the host checked): refused, outputs untouched
    ab, noise_count, total_count, its = run_noalloc(torch, sub, nnz=sub.nnz_path - 3)
    assert its == 0xffffffff and np.all(ab == -1.0) and noise_count == -1.0 and total_count == -1.0

    # path indices of a read class out of order
    bad = copy.copy(sub)
    bad.col = sub.col.copy()
    lengths = np.diff(sub.row_ptr.astype(np.int64))
    r = int(np.argmax(lengths >= 2))
    e = int(sub.row_ptr[r] - sub.row_ptr[0])
    bad.col[e], bad.col[e + 1] = bad.col[e + 1], bad.col[e]
    ab, noise_count, total_count, its = run_noalloc(torch, bad)
    assert its == 0xffffffff and np.all(ab == -1.0)

    # the same cluster untouched still runs
    ab, noise_count, total_count, its = run_noalloc(torch, sub)
    assert 0 < its <= 10000 and np.all(ab >= 0)


def test_noalloc_entry_point_limit_does_not_wrap():
    params = capi.default_em_params()
    status, message = capi.em_cluster_noalloc(0, 0, 10, 4, 1 << 61, 8, 8, 8, 8, 8, params, 8, 8, 8, 8)   # never launched
    assert status == -4 and "fixed area holds" in message
