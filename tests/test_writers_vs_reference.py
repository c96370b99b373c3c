"""SURVEY §8(f) rows 2 and 4 pinned against rpvg's OWN writers (rpvg:src/threaded_output_writer.cpp, compiled unmodified
into oracle/_ref behind a zlib stand-in for htslib's BGZF):

* the `--write-probs` dump written by ProbabilityClusterWriter::addCluster is parsed by this repo's reader, and this repo's
  writer produces the same bytes;
* the three estimate tables (AbundanceEstimatesWriter, HaplotypeAbundanceEstimatesWriter,
  JointHaplotypeAbundanceEstimatesWriter) are written by the reference and their columns compared with the device
  reducers rpvg_b200_tpm, rpvg_b200_haplotype_marginals and rpvg_b200_joint_haplotype_estimates at the 8 significant
  digits the tables hold.
"""

import gzip

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import probs_io, synth


@pytest.fixture(scope="module")
def ref():
    if not loader.reference_available():
        pytest.skip("oracle/_ref is not built")
    return loader.reference()


def path_table(rng, n_paths):
    lengths = rng.integers(200, 5000, size=n_paths).astype(np.uint32)
    eff = lengths - rng.uniform(0, 150, size=n_paths)
    eff[rng.random(n_paths) < 0.03] = 0.0
    return lengths, eff


def names_of(arena):
    return [f"c{c}_p{p}" for c in range(arena.n_clusters) for p in range(int(arena.path_offset[c + 1] - arena.path_offset[c]))]


# ---- f2: the --write-probs dump -----------------------------------------------------------------------------------------

def test_reader_parses_the_reference_writers_dump(ref, tmp_path):
    arena = synth.transcripts_arena(seed=31, n_paths=600, n_reads=4000, n_max=80)
    # values with long decimal expansions, a noise-only read class and equal probabilities inside a read class
    arena.prob[:] = np.round(arena.prob, 12)
    r = int(np.argmax(np.diff(arena.row_ptr.astype(np.int64)) >= 3))
    e = int(arena.row_ptr[r])
    arena.prob[e + 1] = arena.prob[e]
    rng = np.random.default_rng(4)
    lengths, eff = path_table(rng, arena.n_paths)

    dump = loader.ref_write_probs(ref, str(tmp_path / "ref_probs"), arena, lengths, eff)
    back, names, lengths2, eff2 = probs_io.read_probs(dump)

    kept = [c for c in range(arena.n_clusters) if arena.read_offset[c + 1] > arena.read_offset[c]]   # the reference skips empty clusters
    want = arena.select(kept)
    assert back.n_clusters == want.n_clusters
    assert np.array_equal(back.read_offset, want.read_offset) and np.array_equal(back.path_offset, want.path_offset)
    assert np.array_equal(back.row_ptr, want.row_ptr) and np.array_equal(back.col, want.col) and np.array_equal(back.count, want.count)
    assert np.allclose(back.prob, want.prob, rtol=1e-7, atol=0) and np.allclose(back.noise, want.noise, rtol=1e-7, atol=0)   # 8 digits
    all_names = names_of(arena)
    kept_paths = np.concatenate([np.arange(arena.path_offset[c], arena.path_offset[c + 1]) for c in kept]).astype(int)
    assert names == [all_names[i] for i in kept_paths]
    assert np.array_equal(lengths2, lengths[kept_paths]) and np.allclose(eff2, eff[kept_paths], rtol=1e-7)

    # and this repo's writer produces the reference writer's bytes
    ours = str(tmp_path / "our_probs.txt.gz")
    probs_io.write_probs(ours, arena, all_names, lengths, eff, prob_precision=1e-8)
    with gzip.open(dump, "rb") as a, gzip.open(ours, "rb") as b:
        assert a.read() == b.read()


# ---- f4: the estimate tables --------------------------------------------------------------------------------------------

def table(path):
    with open(path) as fh:
        rows = [line.rstrip("\n").split("\t") for line in fh]
    return rows[0], rows[1:]


def close8(got, want):
    """Equal at the 8 significant digits the reference prints."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    return np.allclose(got, want, rtol=6e-8, atol=1e-300)


def random_estimates(rng, n_clusters, ploidy, single_path_sets=False):
    clusters = []
    for c in range(n_clusters):
        n_paths = int(rng.integers(1, 30))
        lengths, eff = path_table(rng, n_paths)
        if single_path_sets:
            sets = [(p,) for p in range(n_paths)]
        else:
            sets = [tuple(sorted(int(x) for x in rng.integers(0, n_paths, size=int(rng.integers(1, ploidy + 1))))) for _ in range(int(rng.integers(0, 40)))]
        posteriors = rng.uniform(0, 1, size=len(sets))
        posteriors[rng.random(len(sets)) < 0.15] = 0.0
        if single_path_sets:
            posteriors[:] = 0.0                                  # the abundance classes leave them at zero
        n_ab = sum(len(s) for s in sets)
        abundances = rng.gamma(0.5, 50.0, size=n_ab)
        abundances[rng.random(n_ab) < 0.2] = 0.0
        noise = float(rng.uniform(0, 5))
        clusters.append(dict(cluster_id=c + 1, length=lengths, effective_length=eff, sets=sets, posteriors=posteriors, abundances=abundances,
                             noise_count=noise, total_count=float(abundances.sum() + noise)))
    return clusters


def flatten(clusters):
    path_offset, set_offset, member_offset, members = [0], [0], [0], []
    for cl in clusters:
        path_offset.append(path_offset[-1] + len(cl["length"]))
        for s in cl["sets"]:
            members.extend(s)
            member_offset.append(len(members))
        set_offset.append(len(member_offset) - 1)
    cat = lambda key: np.concatenate([np.asarray(cl[key], dtype=np.float64) for cl in clusters])
    return (np.array(path_offset, dtype=np.uint64), np.array(set_offset, dtype=np.uint64), np.array(member_offset, dtype=np.uint64),
            np.array(members, dtype=np.uint32), cat("posteriors"), cat("abundances"), cat("effective_length"))


def total_transcript_count(clusters):
    """rpvg:src/main.cpp:1029-1057, in its order."""
    total = 0.0
    for cl in clusters:
        k = 0
        for s in cl["sets"]:
            for p in s:
                if cl["effective_length"][p] > 0:
                    total += cl["abundances"][k] / cl["effective_length"][p]
                k += 1
    return total


@pytest.mark.gpu
def test_tpm_matches_the_reference_abundance_table(ctx, ref, tmp_path):
    rng = np.random.default_rng(21)
    clusters = random_estimates(rng, 60, 1, single_path_sets=True)
    total = total_transcript_count(clusters)
    header, rows = table(loader.ref_write_estimates(ref, 0, str(tmp_path / "abundance"), clusters, total_transcript_count=total, unaligned_read_count=17))
    assert header == ["Name", "ClusterID", "Length", "EffectiveLength", "ReadCount", "TPM"]
    assert rows[-1][0] == "Unknown" and close8(float(rows[-1][4]), sum(cl["noise_count"] for cl in clusters) + 17)

    _, _, _, _, _, abundances, eff = flatten(clusters)
    tpm, dev_total = ctx.tpm(abundances, eff)
    assert abs(dev_total - total) <= 1e-12 * total
    assert len(rows) - 1 == len(tpm)
    assert close8(tpm, [float(r[5]) for r in rows[:-1]])
    assert close8(abundances, [float(r[4]) for r in rows[:-1]])


@pytest.mark.gpu
@pytest.mark.parametrize("ploidy", [2, 3])
def test_marginals_match_the_reference_haplotype_table(ctx, ref, tmp_path, ploidy):
    rng = np.random.default_rng(30 + ploidy)
    clusters = random_estimates(rng, 80, ploidy)
    total = total_transcript_count(clusters)
    header, rows = table(loader.ref_write_estimates(ref, 1, str(tmp_path / "haplotypes"), clusters, ploidy=ploidy, total_transcript_count=total))
    assert header == ["Name", "ClusterID", "Length", "EffectiveLength", "HaplotypeProbability", "ReadCount", "TPM"]

    path_offset, set_offset, member_offset, members, posteriors, abundances, eff = flatten(clusters)
    probs, counts = ctx.haplotype_marginals(path_offset, set_offset, member_offset, members, posteriors, abundances)
    assert len(rows) - 1 == len(probs)
    assert close8(probs, [float(r[4]) for r in rows[:-1]])
    assert close8(counts, [float(r[5]) for r in rows[:-1]])
    per_million = np.where(eff > 0, counts / np.where(eff > 0, eff, 1.0), 0.0) / total * 1e6
    assert close8(per_million, [float(r[6]) for r in rows[:-1]])


@pytest.mark.gpu
@pytest.mark.parametrize("ploidy,min_posterior", [(2, 1e-8), (3, 1e-8), (2, 0.0)])
def test_joint_estimates_match_the_reference_joint_table(ctx, ref, tmp_path, ploidy, min_posterior):
    rng = np.random.default_rng(40 + ploidy)
    clusters = random_estimates(rng, 80, ploidy)
    if min_posterior > 0:
        # the reference asserts that its abundance iterator ends at the end of the list: an estimator never hands over a
        # set below the threshold (rpvg:src/path_abundance_estimator.cpp:576), so the abundances cover the written sets only
        for cl in clusters:
            n_written = sum(len(s) for s, p in zip(cl["sets"], cl["posteriors"]) if p >= min_posterior)
            cl["abundances"] = cl["abundances"][:n_written]
            cl["total_count"] = float(cl["abundances"].sum() + cl["noise_count"])
    total = sum(sum(cl["abundances"]) for cl in clusters) / 1000.0
    header, rows = table(loader.ref_write_estimates(ref, 2, str(tmp_path / "joint"), clusters, ploidy=ploidy, min_posterior=min_posterior,
                                                    total_transcript_count=total, unaligned_read_count=9))
    assert header[:ploidy] == [f"Name_{i + 1}" for i in range(ploidy)]

    # the device call takes the abundances parallel to the members of each cluster's block: pad the tail of every block
    path_offset, set_offset, member_offset, members, posteriors, _, eff = flatten(clusters)
    blocks = []
    for cl in clusters:
        n_members = sum(len(s) for s in cl["sets"])
        blocks.append(np.concatenate([cl["abundances"], np.zeros(n_members - len(cl["abundances"]))]))
    abundances = np.concatenate(blocks)
    noise = [cl["noise_count"] for cl in clusters]

    written, counts, per_million, noise_counts = ctx.joint_haplotype_estimates(path_offset, set_offset, member_offset, members, posteriors, abundances, eff, noise,
                                                                             ploidy, min_posterior, total)
    assert np.array_equal(written, posteriors >= min_posterior)

    body = rows[:-1]
    assert len(body) == int(written.sum())
    k = 0
    for i in np.nonzero(written)[0]:
        row = body[k]
        k += 1
        e0, e1 = int(member_offset[i]), int(member_offset[i + 1])
        c = int(np.searchsorted(set_offset, i, side="right") - 1)
        assert row[:e1 - e0] == [f"c{c + 1}_p{int(m)}" for m in members[e0:e1]] and row[e1 - e0:ploidy] == ["."] * (ploidy - (e1 - e0))
        assert int(row[ploidy]) == c + 1 and close8(posteriors[i], float(row[ploidy + 1]))
        values = [float(x) for x in row[ploidy + 2:]]
        assert close8(counts[e0:e1], values[0:2 * (e1 - e0):2]) and close8(per_million[e0:e1], values[1:2 * (e1 - e0):2])
        assert all(v == 0 for v in values[2 * (e1 - e0):])

    unknown = rows[-1]
    assert unknown[:ploidy] == ["Unknown"] * ploidy
    assert close8(noise_counts + np.float32(9) / ploidy, [float(x) for x in unknown[ploidy + 2::2]])
