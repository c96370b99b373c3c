"""Synthetic boundary inputs for the configurations of BASELINE.json.

The example dataset's binary inputs are not in the reference archive, so every configuration is a seeded synthetic
of the named shape. A generator returns a :class:`ClusterArena` — what ``ReadPathProbabilities::addPathProbs`` and the
pre-boundary merge (rpvg:src/main.cpp:889-973) would hand to ``PathEstimator::estimate`` for each cluster.
"""

from __future__ import annotations

import numpy as np

from .arena import ClusterArena

BASE_SEED = 20240601

# mapq mix of SURVEY.md §8d: noise_prob = max(1e-4, 10^(-mapq/10))
_MAPQ = np.array([60, 30, 10, 1])
_MAPQ_P = np.array([0.85, 0.08, 0.05, 0.02])


def _noise_probs(rng, n):
    mapq = rng.choice(_MAPQ, size=n, p=_MAPQ_P)
    return np.maximum(1e-4, 10.0 ** (-mapq / 10.0))


def power_law_sizes(rng, total, n_max, alpha=1.8):
    """Cluster sizes N_c ~ power-law(alpha) truncated to [1, n_max], drawn until they sum to ``total``."""
    sizes = []
    remaining = total
    e = 1.0 - alpha
    while remaining > 0:
        u = rng.random(max(1024, remaining // 16))
        x = (1.0 - u * (1.0 - float(n_max) ** e)) ** (1.0 / e)
        for n in np.minimum(np.floor(x).astype(np.int64), n_max):
            n = int(min(n, remaining))
            sizes.append(n)
            remaining -= n
            if remaining == 0:
                break
    return np.array(sizes, dtype=np.int64)


def transcripts_arena(seed=BASE_SEED + 2, n_paths=200_000, n_reads=1_000_000, n_max=5000, geom_p=0.25,
                      min_reads_per_cluster=1) -> ClusterArena:
    """Config 2 — "transcripts model EM on synthetic clusters: 1M read classes, 200k paths, power-law cluster
    sizes (1-5,000 paths)". N_c ~ power-law(1.8) on [1, n_max]; R_c = min(1e5, ceil(5 N_c^1.1)) rescaled to sum to
    ``n_reads``; per read class 1 + Geom(geom_p) distinct paths inside a window around a random centre; values
    Dirichlet(1) * (1 - noise); counts 1 + Geom(0.3)."""
    rng = np.random.default_rng(seed)

    sizes = power_law_sizes(rng, n_paths, n_max)
    raw_reads = np.minimum(1e5, np.ceil(5.0 * sizes.astype(np.float64) ** 1.1))
    reads = np.maximum(min_reads_per_cluster, np.round(raw_reads * (n_reads / raw_reads.sum()))).astype(np.int64)

    C = len(sizes)
    read_offset = np.zeros(C + 1, dtype=np.uint64)
    path_offset = np.zeros(C + 1, dtype=np.uint64)
    read_offset[1:] = np.cumsum(reads)
    path_offset[1:] = np.cumsum(sizes)
    R = int(read_offset[-1])

    row_cluster = np.repeat(np.arange(C), reads)
    row_n = sizes[row_cluster]

    k = np.minimum(1 + rng.geometric(geom_p, size=R), row_n)          # distinct paths per read class
    window = np.minimum(row_n, np.maximum(4 * k, 8))
    centre = (rng.random(R) * row_n).astype(np.int64)
    start = np.clip(centre - window // 2, 0, row_n - window)

    row_ptr = np.zeros(R + 1, dtype=np.uint64)
    row_ptr[1:] = np.cumsum(k)
    nnz = int(row_ptr[-1])

    col = np.empty(nnz, dtype=np.uint32)
    prob = np.empty(nnz, dtype=np.float64)
    noise = _noise_probs(rng, R)

    # k distinct sorted positions in [0, window): sorted draws from [0, window - k + 1) plus 0..k-1
    for kk in np.unique(k):
        rows = np.nonzero(k == kk)[0]
        high = (window[rows] - kk + 1)[:, None]
        pos = np.sort(rng.integers(0, high, size=(len(rows), kk)), axis=1) + np.arange(kk)[None, :] + start[rows, None]
        vals = rng.exponential(size=(len(rows), kk))
        vals = vals / vals.sum(axis=1, keepdims=True) * (1.0 - noise[rows, None])
        dst = (row_ptr[rows].astype(np.int64)[:, None] + np.arange(kk)[None, :]).ravel()
        col[dst] = pos.ravel().astype(np.uint32)
        prob[dst] = vals.ravel()

    count = rng.geometric(0.3, size=R).astype(np.uint32)                # 1 + Geom(0.3) on {0,1,..}

    return ClusterArena(read_offset, path_offset, row_ptr, col, prob, noise, count)


def haplotype_cluster(rng, n_transcripts=3, paths_per_transcript=(2, 12), n_haplotypes=40, n_reads=400,
                      window=3, expressed=2) -> ClusterArena:
    """One cluster in the spirit of config 1: transcripts of a gene, each with haplotype-specific
    paths (allele vectors over ~log2(P)+2 variant sites); each haplotype id is assigned one path per transcript; the
    sample is a diplotype (``expressed`` haplotypes); a read covers a window of <= ``window`` sites of one transcript
    and is compatible with every path that matches the true path there (prob ~ 1 / effective length; mismatching
    paths down-weighted by exp(-1.3833 * 5 * mismatches) and dropped below 1e-8)."""
    group_id, alleles, eff_len, path_haps = [], [], [], []
    for t in range(n_transcripts):
        P = int(rng.integers(paths_per_transcript[0], paths_per_transcript[1] + 1))
        sites = int(np.log2(P)) + 2
        seen = set()
        vecs = []
        while len(vecs) < P:
            v = tuple(rng.integers(0, 2, size=sites))
            if v not in seen:
                seen.add(v)
                vecs.append(np.array(v))
        length = float(rng.integers(400, 4000))
        # Zipf-like assignment of haplotype ids to this transcript's paths
        w = 1.0 / np.arange(1, P + 1) ** 1.2
        assign = rng.choice(P, size=n_haplotypes, p=w / w.sum())
        for p in range(P):
            group_id.append(t)
            alleles.append(vecs[p])
            eff_len.append(length)
            path_haps.append(np.nonzero(assign == p)[0])

    n_paths = len(group_id)
    group_id = np.array(group_id, dtype=np.uint32)

    # paths nobody carries get one private haplotype id so that source_count > 0
    next_hap = n_haplotypes
    for p in range(n_paths):
        if len(path_haps[p]) == 0:
            path_haps[p] = np.array([next_hap])
            next_hap += 1

    sample_haps = rng.choice(n_haplotypes, size=expressed, replace=False)
    hap_to_path = {}
    for p in range(n_paths):
        for h in path_haps[p]:
            hap_to_path[(int(h), int(group_id[p]))] = p

    expr = rng.lognormal(sigma=2.0, size=n_transcripts)
    expr /= expr.sum()

    rows = {}
    noise_draws = _noise_probs(rng, n_reads)
    for r in range(n_reads):
        t = int(rng.choice(n_transcripts, p=expr))
        h = int(rng.choice(sample_haps))
        true_path = hap_to_path[(h, t)]
        cand = np.nonzero(group_id == t)[0]
        sites = len(alleles[true_path])
        w = int(rng.integers(1, min(window, sites) + 1))
        s0 = int(rng.integers(0, sites - w + 1))
        noise = float(noise_draws[r])
        lik = {}
        for p in cand:
            mism = int(np.sum(alleles[p][s0:s0 + w] != alleles[true_path][s0:s0 + w]))
            lik[int(p)] = np.exp(-1.3833 * 5 * mism) / eff_len[p]
        tot = sum(lik.values())
        entries = tuple(sorted((p, round(v / tot * (1 - noise), 12)) for p, v in lik.items() if v / tot * (1 - noise) >= 1e-8))
        key = (noise, entries)
        rows[key] = rows.get(key, 0) + 1

    keys = sorted(rows.keys())
    row_ptr, col, prob, noise_arr, count = [0], [], [], [], []
    for (noise, entries) in keys:
        for p, v in entries:
            col.append(p)
            prob.append(v)
        row_ptr.append(len(col))
        noise_arr.append(noise)
        count.append(rows[(noise, entries)])

    source_ptr = np.zeros(n_paths + 1, dtype=np.uint64)
    source_ptr[1:] = np.cumsum([len(h) for h in path_haps])
    return ClusterArena(
        read_offset=np.array([0, len(keys)], dtype=np.uint64), path_offset=np.array([0, n_paths], dtype=np.uint64),
        row_ptr=np.array(row_ptr, dtype=np.uint64), col=np.array(col, dtype=np.uint32), prob=np.array(prob, dtype=np.float64),
        noise=np.array(noise_arr, dtype=np.float64), count=np.array(count, dtype=np.uint32),
        group_id=group_id, source_count=np.array([len(h) for h in path_haps], dtype=np.uint32),
        source_ptr=source_ptr, source_ids=np.concatenate(path_haps).astype(np.uint32))
