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
                      min_reads_per_cluster=1, sizes=None, reads_cap=1e5) -> ClusterArena:
    """Config 2 — "transcripts model EM on synthetic clusters: 1M read classes, 200k paths, power-law cluster
    sizes (1-5,000 paths)". N_c ~ power-law(1.8) on [1, n_max]; R_c = min(1e5, ceil(5 N_c^1.1)) rescaled to sum to
    ``n_reads``; per read class 1 + Geom(geom_p) distinct paths inside a window around a random centre; values
    Dirichlet(1) * (1 - noise); counts 1 + Geom(0.3)."""
    rng = np.random.default_rng(seed)

    # (``sizes``: the paths of every cluster given outright instead of drawn; ``reads_cap``: the cap on R_c before rescaling)
    sizes = power_law_sizes(rng, n_paths, n_max) if sizes is None else np.asarray(sizes, dtype=np.int64)
    raw_reads = np.minimum(reads_cap, np.ceil(5.0 * sizes.astype(np.float64) ** 1.1))
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


# ---- haplotype-specific transcript structures (configs 1, 3 and 4 of BASELINE.json) -------------------------

class PathStructure:
    """Paths grouped by transcript (paths of a transcript are contiguous), each with the haplotype ids that carry it."""

    def __init__(self, path_transcript, path_length, hap_ptr, hap_ids, n_haplotypes):
        self.path_transcript = np.ascontiguousarray(path_transcript, dtype=np.int64)
        self.path_length = np.ascontiguousarray(path_length, dtype=np.int64)
        self.hap_ptr = np.ascontiguousarray(hap_ptr, dtype=np.int64)
        self.hap_ids = np.ascontiguousarray(hap_ids, dtype=np.int64)
        self.n_haplotypes = int(n_haplotypes)
        self.n_transcripts = int(self.path_transcript[-1]) + 1 if len(self.path_transcript) else 0
        self.transcript_ptr = np.searchsorted(self.path_transcript, np.arange(self.n_transcripts + 1))

    @property
    def n_paths(self):
        return len(self.path_transcript)


def example_structure() -> PathStructure:
    """Config 1: the structure of the reference's example pantranscriptome (36,120 paths, 2,177 transcripts, 1,729 haplotype
    ids), frozen by tests/golden/make_c1_structure.py (the reference tree itself does not travel to the GPU box)."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "c1_structure.npz"))
    return PathStructure(z["path_transcript"], z["path_length"], z["hap_ptr"], z["hap_ids"], int(z["hap_ids"].max()) + 1)


def synthetic_structure(rng, n_transcripts, n_haplotypes=5008, paths_per_transcript=None, zipf=1.2, p_min=1, p_max=None) -> PathStructure:
    """Configs 3 and 4: ``n_haplotypes`` haplotype ids; per transcript P paths (drawn from ``paths_per_transcript``, an array
    of counts to sample from — default: the config-1 distribution), every haplotype id assigned to one of them with
    Zipf(``zipf``) weights; paths nobody carries are dropped."""
    if paths_per_transcript is None:
        ex = example_structure()
        paths_per_transcript = np.diff(ex.transcript_ptr)
    sizes = rng.choice(np.asarray(paths_per_transcript), size=n_transcripts)
    if p_max is not None:
        sizes = np.minimum(sizes, p_max)
    sizes = np.maximum(sizes, p_min)
    path_transcript, path_length, hap_ptr, hap_ids = [], [], [0], []
    for t, P in enumerate(sizes):
        w = 1.0 / np.arange(1, P + 1) ** zipf
        assign = rng.choice(P, size=n_haplotypes, p=w / w.sum())
        order = np.argsort(assign, kind="stable")
        counts = np.bincount(assign, minlength=P)
        length = int(rng.integers(400, 4000))
        start = 0
        for p in range(P):
            if counts[p] == 0:
                continue
            path_transcript.append(t)
            path_length.append(length)
            hap_ids.append(order[start:start + counts[p]])
            start += counts[p]
            hap_ptr.append(hap_ptr[-1] + counts[p])
        if start != n_haplotypes:   # (paths with no carrier were skipped without consuming ids: cannot happen)
            raise AssertionError
    return PathStructure(path_transcript, path_length, hap_ptr, np.concatenate(hap_ids), n_haplotypes)


def _allele_vectors(rng, P):
    """P distinct allele vectors over ~log2(P) + 2 variant sites."""
    sites = int(np.log2(max(P, 1))) + 2
    codes = rng.choice(1 << sites, size=P, replace=False)
    return ((codes[:, None] >> np.arange(sites)[None, :]) & 1).astype(np.int8)


def haplotype_arena(structure: PathStructure, n_reads, seed, transcripts=None, gene_geom=0.35, ploidy=2, window=3, cross_transcript=0.25,
                    frag_mean=250.0, frag_sd=50.0, single_transcript_clusters=False) -> ClusterArena:
    """Read classes over haplotype-specific transcript paths (SURVEY.md §8d, C1 model). Transcripts are grouped into
    clusters ("genes") of 1 + Geom(``gene_geom``) transcripts; the sample carries ``ploidy`` random haplotype ids;
    expression is log-normal(sigma = 2) over the transcripts; a read class = (transcript, true path, window of <= ``window``
    variant sites, mapping quality, fragment length) with a Poisson count; it is compatible with every path of the transcript —
    probability ~ exp(-1.3833 * 5 * mismatches on the window) * N(fragment length + the path's indel shift; frag_mean,
    frag_sd) / effective length, normalised, times (1 - noise), entries below 1e-8
    folded into the noise like ReadPathProbabilities::addPathProbs does — and, with probability ``cross_transcript``, with
    the paths of a second transcript of its cluster at one extra mismatch. Identical rows of a cluster are merged.
    The arena carries PathInfo's group_id (transcript), source_ids (haplotype ids) and source_count."""
    rng = np.random.default_rng(seed)
    st = structure
    transcripts = np.arange(st.n_transcripts) if transcripts is None else np.asarray(transcripts)

    # clusters of transcripts
    genes, k = [], 0
    while k < len(transcripts):
        size = 1 if single_transcript_clusters else int(rng.geometric(1.0 - gene_geom))   # 1 + Geom(0.35) on {0, 1, ..}
        genes.append(transcripts[k:k + size])
        k += size

    sample = rng.choice(st.n_haplotypes, size=ploidy, replace=False)

    # the sample's path of every transcript (transcripts none of its haplotypes covers give no reads)
    carried = np.zeros(st.n_paths, dtype=bool)
    for h in sample:
        carried |= np.bincount(np.repeat(np.arange(st.n_paths), np.diff(st.hap_ptr))[st.hap_ids == h], minlength=st.n_paths) > 0
    covered = np.array([carried[st.transcript_ptr[t]:st.transcript_ptr[t + 1]].any() for t in transcripts])

    expr = rng.lognormal(sigma=2.0, size=len(transcripts)) * covered
    expr = expr / expr.sum() * n_reads
    expr_of = dict(zip(transcripts.tolist(), expr.tolist()))
    noise_levels = np.maximum(1e-4, 10.0 ** (-_MAPQ / 10.0))

    read_offset, path_offset, row_ptr = [0], [0], [0]
    col, prob, noise, count = [], [], [], []
    group_id, source_count, source_ptr, source_ids = [], [], [0], []

    for gene in genes:

        infos = []   # per transcript of the cluster: (first local path, alleles, effective lengths, expressed local paths, indel shifts)
        n_local = 0
        for t in gene:
            p0, p1 = int(st.transcript_ptr[t]), int(st.transcript_ptr[t + 1])
            P = p1 - p0
            alleles = _allele_vectors(rng, P)
            eff = np.maximum(1.0, st.path_length[p0:p1] - frag_mean).astype(np.float64)
            expressed = []
            for h in sample:   # one path per haplotype (a path two of them share is listed twice: homozygous)
                for p in np.nonzero(carried[p0:p1])[0]:
                    ids = st.hap_ids[st.hap_ptr[p0 + p]:st.hap_ptr[p0 + p + 1]]
                    if ids[np.searchsorted(ids, h) % len(ids)] == h:
                        expressed.append(int(p))
                        break
            infos.append((n_local, alleles, eff, expressed, rng.integers(-3, 4, size=P).astype(np.float64)))
            for p in range(P):
                ids = st.hap_ids[st.hap_ptr[p0 + p]:st.hap_ptr[p0 + p + 1]]
                group_id.append(int(t))
                source_count.append(len(ids))
                source_ids.append(ids)
                source_ptr.append(source_ptr[-1] + len(ids))
            n_local += P

        rows = {}
        for ti, t in enumerate(gene):
            base, alleles, eff, expressed, shift = infos[ti]
            if not expressed:
                continue
            sites = alleles.shape[1]
            windows = [(s0, w) for w in range(1, min(window, sites) + 1) for s0 in range(sites - w + 1)]
            lam = expr_of[int(t)] / (len(expressed) * len(windows))
            for p_true in expressed:
                for (s0, w) in windows:
                    n = int(rng.poisson(lam))
                    if n == 0:
                        continue
                    # the reads of this (true path, window): mapping quality and fragment length of each
                    level_idx = rng.choice(len(noise_levels), size=n, p=_MAPQ_P)
                    frag = np.clip(np.round(rng.normal(frag_mean, frag_sd, size=n)), 50, 3 * frag_mean).astype(np.int64)
                    keys, cnts = np.unique(level_idx * 100000 + frag, return_counts=True)

                    mism = (alleles[:, s0:s0 + w] != alleles[p_true, s0:s0 + w]).sum(axis=1)
                    lik = np.exp(-1.3833 * 5 * mism) / eff
                    cols = base + np.arange(len(lik))
                    shifts = shift
                    if len(gene) > 1 and rng.random() < cross_transcript:
                        t2 = int(rng.integers(0, len(gene) - 1))
                        t2 += t2 >= ti
                        base2, alleles2, eff2, _, shift2 = infos[t2]
                        sites2 = alleles2.shape[1]
                        w2 = min(w, sites2)
                        s2 = int(rng.integers(0, sites2 - w2 + 1))
                        ref2 = int(rng.integers(0, len(eff2)))
                        mism2 = (alleles2[:, s2:s2 + w2] != alleles2[ref2, s2:s2 + w2]).sum(axis=1) + 1
                        lik = np.concatenate([lik, np.exp(-1.3833 * 5 * mism2) / eff2])
                        cols = np.concatenate([cols, base2 + np.arange(len(eff2))])
                        shifts = np.concatenate([shift, shift2])
                        order = np.argsort(cols, kind="stable")
                        cols, lik, shifts = cols[order], lik[order], shifts[order]

                    # fragment length on each path (paths differ by small indels): one row per distinct (mapq, length)
                    f = (keys % 100000).astype(np.float64)
                    z = (f[:, None] + shifts[None, :] - frag_mean) / frag_sd
                    lik2 = lik[None, :] * np.exp(-0.5 * z * z)
                    share = lik2 / lik2.sum(axis=1, keepdims=True)
                    levels = noise_levels[keys // 100000]
                    keep = share >= 1e-8
                    nz = levels + (1 - levels) * np.where(keep, 0.0, share).sum(axis=1)
                    values = np.round(share * (1 - levels)[:, None], 14)

                    for k in range(len(keys)):
                        m = keep[k]
                        key = (float(nz[k]), cols[m].tobytes(), values[k, m].tobytes())
                        rows[key] = rows.get(key, 0) + int(cnts[k])

        for key in sorted(rows):
            nz, cbytes, vbytes = key
            c_arr = np.frombuffer(cbytes, dtype=np.int64)
            col.append(c_arr.astype(np.uint32))
            prob.append(np.frombuffer(vbytes, dtype=np.float64))
            row_ptr.append(row_ptr[-1] + len(c_arr))
            noise.append(nz)
            count.append(rows[key])

        read_offset.append(len(noise))
        path_offset.append(path_offset[-1] + n_local)

    cat = lambda parts, dtype: np.concatenate(parts).astype(dtype) if parts else np.zeros(0, dtype=dtype)
    return ClusterArena(
        read_offset=np.array(read_offset, dtype=np.uint64), path_offset=np.array(path_offset, dtype=np.uint64), row_ptr=np.array(row_ptr, dtype=np.uint64),
        col=cat(col, np.uint32), prob=cat(prob, np.float64), noise=np.array(noise, dtype=np.float64), count=np.array(count, dtype=np.uint32),
        group_id=np.array(group_id, dtype=np.uint32), source_count=np.array(source_count, dtype=np.uint32),
        source_ptr=np.array(source_ptr, dtype=np.uint64), source_ids=cat(source_ids, np.uint32))


def example_arena(n_reads=100_000, seed=BASE_SEED + 1, transcripts=None) -> ClusterArena:
    """Config 1 — "example/ dataset: 100k read pairs, 36,120 haplotype-specific transcripts, haplotype-transcripts model"."""
    return haplotype_arena(example_structure(), n_reads, seed, transcripts=transcripts)


def thousand_genomes_arena(n_transcripts=2000, reads_per_transcript=300, seed=BASE_SEED + 3, n_haplotypes=5008) -> ClusterArena:
    """Config 3 — haplotype-transcripts at 1000GP scale: 5,008 haplotype ids, paths per transcript from the config-1
    distribution, Zipf(1.2) haplotype assignment. The full configuration has 100,000 transcripts and 30 M read pairs;
    ``n_transcripts`` and ``reads_per_transcript`` are the knobs (the structure of every transcript is independent of them)."""
    rng = np.random.default_rng(seed)
    st = synthetic_structure(rng, n_transcripts, n_haplotypes=n_haplotypes)
    return haplotype_arena(st, n_transcripts * reads_per_transcript, seed + 1000)


def haplotypes_arena(n_clusters=200, n_max=2000, reads_per_path=40, seed=BASE_SEED + 4, n_haplotypes=5008, ploidy=2, alpha=1.3) -> ClusterArena:
    """Config 4 — haplotypes model: one transcript per cluster with N_c haplotype-specific paths, N_c ~ power-law on
    [2, n_max]; read classes from the config-1 window model. ``ploidy`` 2 for the diploid (bounded) posteriors, 4 for
    --use-hap-gibbs."""
    rng = np.random.default_rng(seed)
    u = rng.random(n_clusters)
    e = 1.0 - alpha
    sizes = np.floor((2.0 ** e + u * (float(n_max) ** e - 2.0 ** e)) ** (1.0 / e)).astype(np.int64)
    st = synthetic_structure(rng, n_clusters, n_haplotypes=n_haplotypes, paths_per_transcript=sizes, zipf=0.6, p_min=2)
    n_reads = int(reads_per_path * st.n_paths)
    return haplotype_arena(st, n_reads, seed + 1000, ploidy=ploidy, single_transcript_clusters=True, cross_transcript=0.0)


def strains_arena(seed=BASE_SEED + 5, n_paths=200_000, n_reads=1_000_000, n_max=5000, redundant=0.3, jitter=0.05) -> ClusterArena:
    """Config 5 — strains model: config 2's matrices with ``redundant`` of the paths of every cluster turned into near
    duplicates of another path of the cluster (same read classes, values times exp(N(0, jitter)), rows renormalised to
    1 - noise), so that the minimum path cover is not trivial."""
    base = transcripts_arena(seed=seed, n_paths=n_paths, n_reads=n_reads, n_max=n_max)
    rng = np.random.default_rng(seed + 1000)

    P = base.n_paths
    cluster_of_path = np.repeat(np.arange(base.n_clusters), np.diff(base.path_offset.astype(np.int64)))
    sizes = np.diff(base.path_offset.astype(np.int64))
    local = np.arange(P) - np.repeat(base.path_offset[:-1].astype(np.int64), sizes)

    # duplicates: a path may become the copy of an earlier path of its cluster that stays original
    is_dup = (rng.random(P) < redundant) & (local > 0)
    source = np.arange(P)
    originals = np.nonzero(~is_dup)[0]
    dup_ids = np.nonzero(is_dup)[0]
    # the closest original before it within the cluster (local index 0 is never a duplicate)
    pos = np.searchsorted(originals, dup_ids) - 1
    back = rng.integers(0, 8, size=len(dup_ids))
    first_orig = np.searchsorted(originals, base.path_offset[:-1].astype(np.int64))[cluster_of_path[dup_ids]]
    source[dup_ids] = originals[np.maximum(first_orig, pos - back)]

    # copies of every original, as a CSR over original paths
    order = np.argsort(source[dup_ids], kind="stable")
    copy_of = dup_ids[order]
    copy_ptr = np.searchsorted(source[dup_ids][order], np.arange(P + 1))

    R = base.n_reads
    rows = np.repeat(np.arange(R), np.diff(base.row_ptr.astype(np.int64)))
    cluster_of_row = np.repeat(np.arange(base.n_clusters), np.diff(base.read_offset.astype(np.int64)))
    gcol = base.col.astype(np.int64) + base.path_offset[:-1].astype(np.int64)[cluster_of_row[rows]]

    keep = ~is_dup[gcol]                      # entries of paths that became copies disappear ...
    orphan = np.bincount(rows[keep], minlength=R) == 0
    keep |= orphan[rows]                      # ... except in read classes that would be left without any path
    e_rows, e_cols, e_vals = rows[keep], gcol[keep], base.prob[keep]
    n_copies = (copy_ptr[e_cols + 1] - copy_ptr[e_cols])
    rep = np.repeat(np.arange(len(e_rows)), n_copies)
    offs = np.arange(len(rep)) - np.repeat(np.cumsum(n_copies) - n_copies, n_copies)
    c_rows = e_rows[rep]
    c_cols = copy_of[copy_ptr[e_cols[rep]] + offs] if len(rep) else np.zeros(0, dtype=np.int64)
    c_vals = e_vals[rep] * np.exp(rng.normal(0.0, jitter, size=len(rep)))

    all_rows = np.concatenate([e_rows, c_rows])
    all_cols = np.concatenate([e_cols, c_cols])
    all_vals = np.concatenate([e_vals, c_vals])
    order = np.lexsort((all_cols, all_rows))
    all_rows, all_cols, all_vals = all_rows[order], all_cols[order], all_vals[order]

    row_sum = np.bincount(all_rows, weights=all_vals, minlength=R)
    all_vals = all_vals / row_sum[all_rows] * (1.0 - base.noise[all_rows])

    row_ptr = np.zeros(R + 1, dtype=np.uint64)
    row_ptr[1:] = np.cumsum(np.bincount(all_rows, minlength=R))
    local_cols = all_cols - base.path_offset[:-1].astype(np.int64)[cluster_of_row[all_rows]]

    return ClusterArena(base.read_offset, base.path_offset, row_ptr, local_cols.astype(np.uint32), all_vals, base.noise, base.count)
