"""Device sort + merge of the read records of a cluster (rpvg_b200_merge_read_records) against the reference's own
std::sort(operator<) + quickMergeIdentical scan (rpvg:src/main.cpp:953-973), run through oracle/_ref. Bit-exact: which
records are kept, in which order, with which counts."""

import numpy as np
import pytest

from oracle import loader

pytestmark = pytest.mark.gpu


def make_cluster(rng, n_records, n_paths, n_templates):
    """Records drawn from a few templates (same path-id structure), probabilities jittered by exact copies, by less than
    prob_precision (merge) or by more (no merge); noise values from a small set. Groups ascending by probability."""
    templates = []
    for _ in range(n_templates):
        n_groups = int(rng.integers(0, 4))
        groups = []
        for _ in range(n_groups):
            ids = np.sort(rng.choice(n_paths, size=int(rng.integers(1, 4)), replace=False)).astype(np.uint32)
            groups.append(ids)
        base = np.sort(rng.uniform(0.05, 0.9, size=n_groups))
        templates.append((groups, base))
    noise_values = np.array([1e-4, 1e-3, 0.1, 1.0])
    count, noise, entry_ptr, entry_prob, path_ptr, ids = [], [], [0], [], [0], []
    for _ in range(n_records):
        groups, base = templates[int(rng.integers(0, n_templates))]
        kind = rng.integers(0, 3)
        jitter = {0: 0.0, 1: 3e-9, 2: 5e-7}[int(kind)]
        probs = base + jitter * rng.integers(0, 2, size=len(base))
        nz = noise_values[int(rng.integers(0, 4))] + (2e-9 if rng.random() < 0.2 else 0.0)
        count.append(int(rng.integers(1, 6)))
        noise.append(nz)
        for p, g in zip(probs, groups):
            entry_prob.append(p)
            ids.extend(g.tolist())
            path_ptr.append(len(ids))
        entry_ptr.append(len(entry_prob))
    return (np.array(count, dtype=np.uint32), np.array(noise), np.array(entry_ptr, dtype=np.uint64), np.array(entry_prob),
            np.array(path_ptr, dtype=np.uint64), np.array(ids, dtype=np.uint32))


def flatten_kept(cluster, reps, counts):
    count, noise, entry_ptr, entry_prob, path_ptr, ids = cluster
    o_noise, o_count, o_entry_ptr, o_prob, o_path_ptr, o_ids = [], [], [0], [], [0], []
    for r, c in zip(reps, counts):
        o_noise.append(noise[r])
        o_count.append(c)
        for e in range(int(entry_ptr[r]), int(entry_ptr[r + 1])):
            o_prob.append(entry_prob[e])
            o_ids.extend(ids[int(path_ptr[e]):int(path_ptr[e + 1])].tolist())
            o_path_ptr.append(len(o_ids))
        o_entry_ptr.append(len(o_prob))
    return (np.array(o_noise), np.array(o_count, dtype=np.uint32), np.array(o_entry_ptr, dtype=np.uint64), np.array(o_prob),
            np.array(o_path_ptr, dtype=np.uint64), np.array(o_ids, dtype=np.uint32))


@pytest.mark.parametrize("seed, sizes", [(1, [40, 1, 0, 300]), (2, [2000, 7, 513]), (3, [5000])])
def test_merge_matches_reference(ctx, reference, seed, sizes):
    rng = np.random.default_rng(seed)
    clusters = [make_cluster(rng, n, 12, 6) for n in sizes]

    # one batch: concatenate
    read_offset = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    group_ptr, group_prob, id_ptr, ids, noise, count = [np.zeros(1, dtype=np.uint64)], [], [np.zeros(1, dtype=np.uint64)], [], [], []
    g_base = i_base = 0
    for cl in clusters:
        c_count, c_noise, c_entry_ptr, c_prob, c_path_ptr, c_ids = cl
        group_ptr.append(c_entry_ptr[1:] + np.uint64(g_base))
        id_ptr.append(c_path_ptr[1:] + np.uint64(i_base))
        group_prob.append(c_prob); ids.append(c_ids); noise.append(c_noise); count.append(c_count)
        g_base += len(c_prob); i_base += len(c_ids)
    n_merged, record_class, merged_rep, merged_count = ctx.merge_read_records(
        read_offset, np.concatenate(group_ptr), np.concatenate(group_prob) if g_base else np.zeros(0), np.concatenate(id_ptr),
        np.concatenate(ids) if i_base else np.zeros(0, dtype=np.uint32), np.concatenate(noise) if len(noise) else np.zeros(0), np.concatenate(count))

    for c, cl in enumerate(clusters):
        r0 = int(read_offset[c])
        k = int(n_merged[c])
        if sizes[c] == 0:
            assert k == 0
            continue
        mine = flatten_kept(cl, merged_rep[r0:r0 + k], merged_count[r0:r0 + k])
        ref = loader.ref_merge_read_records(reference, 12, *cl)
        assert k == len(ref[0]), (k, len(ref[0]))
        for x, y in zip(mine, ref):
            assert np.array_equal(x, y)
        # every input record is accounted for exactly once
        assert merged_count[r0:r0 + k].sum() == cl[0].sum()
        cls = record_class[r0:r0 + sizes[c]]
        assert cls.max() == k - 1 and np.array_equal(np.bincount(cls, weights=cl[0], minlength=k).astype(np.uint32), merged_count[r0:r0 + k])
