"""Read-class collapse on the device (rpvg_b200_collapse) against the CPU oracle, which is itself bit-identical to the
reference's readCollapseProbabilityMatrix (tests/test_oracle_vs_reference.py). Bar: partition of rows into classes, class
counts and class order exact; representative values equal within the comparator's own tolerance (100 eps relative)."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make_problem(rng, rows, cols, distinct):
    base = rng.random((distinct, cols)) * (rng.random((distinct, cols)) < 0.6)
    base[base.sum(axis=1) == 0, 0] = 0.5
    noise_base = np.maximum(1e-4, 10.0 ** (-rng.choice([60, 30, 10, 1], size=distinct) / 10.0))
    base = base / base.sum(axis=1, keepdims=True) * (1 - noise_base[:, None])
    pick = rng.integers(0, distinct, size=rows)
    mat = base[pick].copy()
    noise = noise_base[pick].copy()
    jitter = rng.integers(0, 5, size=rows)
    mat[jitter == 1] *= (1 + 2.2e-16)      # equal for the comparator, bitwise different
    mat[jitter == 2, 0] += 3e-9            # inside prob_precision, outside the comparator's tolerance
    mat[jitter == 3, cols - 1] += 5e-8     # outside prob_precision
    noise[jitter == 4] += 4e-9
    counts = rng.geometric(0.3, size=rows).astype(np.float64)
    return mat, noise, counts


def to_csr(problems):
    row_offset, row_ptr, col, val, noise, count = [0], [0], [], [], [], []
    for mat, nz, cnt in problems:
        for i in range(mat.shape[0]):
            js = np.nonzero(mat[i])[0]
            col.extend(js.tolist())
            val.extend(mat[i, js].tolist())
            row_ptr.append(len(col))
        noise.extend(nz.tolist())
        count.extend(cnt.tolist())
        row_offset.append(row_offset[-1] + mat.shape[0])
    return (np.array(row_offset, dtype=np.uint64), np.array(row_ptr, dtype=np.uint64), np.array(col, dtype=np.uint32), np.array(val), np.array(noise),
            np.array(count))


def oracle_collapse(port, mat, noise, counts):
    rows, cols = mat.shape
    full = np.asfortranarray(np.hstack([mat, noise[:, None]]))
    out = np.zeros(rows * (cols + 1)); cnt = np.zeros(rows); cls = np.zeros(rows, dtype=np.uint32)
    n = port.oracle_collapse_dense(1e-8, rows, cols + 1, full.ctypes.data, counts.ctypes.data, out.ctypes.data, cnt.ctypes.data, cls.ctypes.data)
    return n, out[:n * (cols + 1)].reshape((n, cols + 1), order="F"), cnt[:n], cls


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_collapse_matches_oracle(ctx, port, seed):
    rng = np.random.default_rng(seed)
    problems = [make_problem(rng, int(rows), int(cols), int(distinct)) for rows, cols, distinct in
                [(1, 1, 1), (40, 3, 7), (300, 6, 40), (2000, 4, 150), (5000, 12, 900), (17, 2, 17)]]
    row_class, n_classes, class_rep, class_count = ctx.collapse(*to_csr(problems))

    r0 = c0 = 0
    for p, (mat, noise, counts) in enumerate(problems):
        rows = mat.shape[0]
        n, out, cnt, cls = oracle_collapse(port, mat, noise, counts)
        assert n_classes[p] == n
        assert n < rows or rows < 20
        assert np.array_equal(row_class[r0:r0 + rows], cls), f"problem {p}: partition / class order differs"
        assert np.array_equal(class_count[c0:c0 + n], cnt)
        reps = class_rep[c0:c0 + n].astype(np.int64) - r0
        assert np.array_equal(cls[reps], np.arange(n)), "a representative must be a member of its class"
        rep_vals = np.hstack([mat[reps], noise[reps][:, None]])
        assert np.allclose(rep_vals, out, rtol=100 * 2.3e-16, atol=0)
        r0 += rows
        c0 += n


def test_collapse_edge_cases(ctx, port):
    """Empty problem, noise-only rows (no entries), all rows identical, all rows distinct."""
    same = (np.tile(np.array([[0.2, 0.0, 0.7]]), (50, 1)), np.full(50, 0.1), np.arange(1, 51, dtype=np.float64))
    noise_only = (np.zeros((5, 2)), np.array([1.0, 1.0, 0.5, 1.0, 0.5]), np.array([1.0, 2.0, 3.0, 4.0, 5.0]))
    rng = np.random.default_rng(9)
    distinct = (rng.random((30, 3)), rng.random(30) * 0.1, np.ones(30))
    empty = (np.zeros((0, 3)), np.zeros(0), np.zeros(0))
    problems = [same, empty, noise_only, distinct]
    row_class, n_classes, class_rep, class_count = ctx.collapse(*to_csr(problems))
    assert list(n_classes) == [1, 0, 2, 30]
    assert class_count[0] == same[2].sum()
    n, out, cnt, cls = oracle_collapse(port, *noise_only)
    assert n == 2 and np.array_equal(class_count[1:3], cnt) and np.array_equal(row_class[50:55], cls)
    n, out, cnt, cls = oracle_collapse(port, *distinct)
    assert np.array_equal(row_class[55:], cls)


def test_collapse_hundred_thousand_rows(ctx, port):
    """One problem of 100,000 rows (a C5-sized cluster: 25k-100k read classes): near duplicates inside the precision,
    ulp-level twins and rows just outside it, so that certain starts and ambiguous stretches both occur; plus a problem
    made of one long chain of rows each within the precision of its predecessor (the worst case for the cut into
    stretches: one thread walks it all)."""
    rng = np.random.default_rng(11)
    big = make_problem(rng, 100_000, 8, 6000)

    chain_rows = 3000
    chain = np.zeros((chain_rows, 3))
    chain[:, 0] = 0.3 + 6e-9 * np.arange(chain_rows)       # 6e-9 < precision between neighbours, drifting far over the chain
    chain[:, 2] = 0.2
    chain_problem = (chain, np.full(chain_rows, 0.01), np.ones(chain_rows))

    problems = [big, chain_problem]
    row_class, n_classes, class_rep, class_count = ctx.collapse(*to_csr(problems))

    r0 = c0 = 0
    for p, (mat, noise, counts) in enumerate(problems):
        rows = mat.shape[0]
        n, out, cnt, cls = oracle_collapse(port, mat, noise, counts)
        assert n_classes[p] == n
        assert np.array_equal(row_class[r0:r0 + rows], cls), f"problem {p}: partition / class order differs"
        assert np.array_equal(class_count[c0:c0 + n], cnt)
        r0 += rows
        c0 += n
    assert n_classes[0] > 5000 and 100 < n_classes[1] < chain_rows
