"""rpvg_b200_add_path_probs against the reference's own ReadPathProbabilities::addPathProbs
(rpvg:src/read_path_probabilities.cpp:74-221, compiled unmodified into oracle/_ref): first the known-answer cases the
reference's own test file holds (rpvg:src/tests/read_path_probabilities_test.cpp:9-205), then seeded random reads —
identical grouping and order, values within 1e-12 relative (device exp/log vs glibc)."""

import numpy as np
import pytest

from oracle import loader
from rpvg_b200 import capi

pytestmark = pytest.mark.gpu

FRAG = dict(frag_mean=10.0, frag_sd=2.0, sd_max_multi=10)
LOWEST_SCORE = -2 ** 31   # numeric_limits<int32_t>::lowest(): the noise alignment path of the reference's tests


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def ref():
    if not loader.reference_available():
        pytest.skip("oracle/_ref is not built")
    return loader.reference()


def frag_log_probs(ref, frag_length):
    return np.array([ref.ref_frag_log_prob(FRAG["frag_mean"], FRAG["frag_sd"], FRAG["sd_max_multi"], int(f)) for f in frag_length], dtype=np.float64)


def run_both(ctx, ref, eff, mapq, reads, min_noise=0.0, precision=1e-8, path_group=None, source_count=None, single_end=False):
    """reads: list of lists of (score_sum, align_length, frag_length, [path ids]); the last entry of a read is its noise
    alignment path (no ids)."""
    align_ptr, score, alen, flen, id_ptr, ids = [0], [], [], [], [0], []
    for read in reads:
        for (s, l, f, p) in read:
            score.append(s); alen.append(l); flen.append(f); ids.extend(p); id_ptr.append(len(ids))
        align_ptr.append(len(score))
    n_groups = 0 if path_group is None else int(max(path_group)) + 1
    expect = loader.ref_add_path_probs(ref, eff, mapq, align_ptr, score, alen, flen, id_ptr, ids, is_single_end=single_end, min_noise_prob=min_noise,
                                       prob_precision=precision, path_group=path_group, source_count=source_count, num_groups=n_groups, **FRAG)
    got = ctx.add_path_probs([0, len(reads)], [0, len(eff)], eff, mapq, align_ptr, score, alen, None if single_end else frag_log_probs(ref, flen), id_ptr, ids,
                             path_group=path_group, source_count=source_count, min_noise_prob=min_noise, prob_precision=precision)
    return expect, got


def assert_same(expect, got, rel=1e-12):
    e_noise, e_gptr, e_prob, e_iptr, e_ids = expect
    g_noise, g_gptr, g_prob, g_iptr, g_ids = got
    assert np.array_equal(e_gptr, g_gptr)
    assert np.array_equal(e_iptr, g_iptr)
    assert np.array_equal(e_ids, g_ids)
    assert np.allclose(g_noise, e_noise, rtol=rel, atol=0)
    assert np.allclose(g_prob, e_prob, rtol=rel, atol=0)


def close(a, b):
    # Utils::doubleCompare (rpvg:src/utils.hpp:87-93)
    return a == b or abs(a - b) < abs(min(a, b)) * 100 * np.finfo(np.float64).eps


def test_reference_known_answers(ctx, ref):
    noise_path = (LOWEST_SCORE, 0, 0, [])

    # read_path_probabilities_test.cpp:9-34
    expect, got = run_both(ctx, ref, [3, 3], [10], [[(3, 5, 10, [0, 1]), noise_path]])
    assert_same(expect, got)
    assert close(got[0][0], 0.1) and len(got[2]) == 1 and close(got[2][0], 0.45) and list(got[4]) == [0, 1]

    # :36-48 improbable fragment length
    expect, got = run_both(ctx, ref, [3, 3], [10], [[(3, 5, 10000, [0, 1]), noise_path]])
    assert_same(expect, got)
    assert abs(got[2][0] - 0.45) < 1e-8

    # :50-78 two alignment paths (path index 3 = rpvg's path id 50)
    eff4 = [3, 3, 3, 3]
    two = [[(3, 5, 10, [0, 1]), (5, 8, 15, [3]), noise_path]]
    expect, got = run_both(ctx, ref, eff4, [10], two)
    assert_same(expect, got)
    assert len(got[2]) == 2 and close(got[2][0], 0.233044027062125) and close(got[2][1], 0.333477986468937)
    assert list(got[4][int(got[3][0]):int(got[3][1])]) == [3] and list(got[4][int(got[3][1]):int(got[3][2])]) == [0, 1]

    # :80-94 precision 0.1 merges all three
    expect, got = run_both(ctx, ref, [3, 3, 3, 2], [10], two, precision=0.1)
    assert_same(expect, got)
    assert len(got[2]) == 1 and close(got[2][0], 0.3) and list(got[4]) == [0, 1, 3]

    # :96-112 the longest alignment path wins
    longest = [[(3, 5, 10, [0, 1]), (5, 8, 15, [3]), (3, 10, 10, [3]), noise_path]]
    expect, got = run_both(ctx, ref, eff4, [10], longest, precision=0.1)
    assert_same(expect, got)
    assert len(got[2]) == 1 and close(got[2][0], 0.3) and list(got[4]) == [0, 1, 3]

    # :114-132 equal length: the highest scoring one
    highest = [[(3, 5, 10, [0, 1]), (5, 8, 15, [3]), (3, 8, 15, [3]), noise_path]]
    expect, got = run_both(ctx, ref, eff4, [10], highest, precision=0.1)
    assert_same(expect, got)
    assert len(got[2]) == 2 and abs(got[2][0] - 0.233044027062125) < 1e-8 and abs(got[2][1] - 0.333477986468937) < 1e-8

    # :135-160 the noise alignment path's score
    score = int(-2.302585 / 1e-6)
    expect, got = run_both(ctx, ref, [3, 3], [10], [[(3, 5, 10, [0, 1]), (score, 0, 0, [])]])
    assert_same(expect, got)
    assert close(got[0][0], 0.190000008369464) and close(got[2][0], 0.404999995815267)
    expect, got = run_both(ctx, ref, [3, 3], [10], [[(3, 5, 10, [0, 1]), (0, 0, 0, [])]])
    assert_same(expect, got)
    assert close(got[0][0], 1.0) and len(got[2]) == 0

    # :162-176 effective lengths
    expect, got = run_both(ctx, ref, [3, 2], [10], [[(3, 5, 10, [0, 1]), noise_path]])
    assert_same(expect, got)
    assert close(got[2][0], 0.36) and close(got[2][1], 0.54)

    # :178-191 base noise probability
    expect, got = run_both(ctx, ref, [3, 3], [10], [[(3, 5, 10, [0, 1]), (int(-5.0 / 1e-6), 0, 0, [])]], min_noise=0.3)
    assert_same(expect, got)
    assert close(got[0][0], 0.304716562899359) and close(got[2][0], 0.347641718550320)


def random_reads(rng, n_paths, n_reads, zero_eff=0.05):
    eff = rng.integers(50, 3000, size=n_paths).astype(np.float64)
    eff[rng.random(n_paths) < zero_eff] = 0.0
    mapq, reads = [], []
    for _ in range(n_reads):
        mapq.append(int(rng.choice([0, 1, 10, 30, 60], p=[0.03, 0.07, 0.1, 0.2, 0.6])))
        read = []
        for _ in range(int(rng.integers(1, 5))):
            k = int(rng.integers(1, min(n_paths, 12) + 1))
            start = int(rng.integers(0, n_paths))
            ids = sorted({int((start + int(x)) % n_paths) for x in rng.integers(0, 20, size=k)})
            rng.shuffle(ids)
            read.append((int(rng.integers(-5, 40)), int(rng.integers(20, 300)), int(rng.integers(1, 40)), list(ids)))
        read.append((int(rng.choice([0, -1000, -200000, -3000000, LOWEST_SCORE])), 0, 0, []))
        reads.append(read)
    # every read needs one path of positive effective length (rpvg asserts otherwise)
    good = int(np.argmax(eff > 0))
    for read in reads:
        if all(eff[p] == 0 for a in read for p in a[3]):
            read[0][3].append(good)
    return eff, mapq, reads


@pytest.mark.parametrize("seed,single_end", [(1, False), (2, False), (3, True)])
def test_random_reads_match_the_reference(ctx, ref, seed, single_end):
    rng = np.random.default_rng(seed)
    eff, mapq, reads = random_reads(rng, n_paths=60, n_reads=800)
    expect, got = run_both(ctx, ref, eff, mapq, reads, min_noise=1e-4, single_end=single_end)
    assert_same(expect, got)
    assert len(got[2]) > 1000


def test_collapse_groups_matches_the_reference(ctx, ref):
    rng = np.random.default_rng(7)
    eff, mapq, reads = random_reads(rng, n_paths=48, n_reads=500)
    path_group = rng.integers(0, 9, size=48).astype(np.uint32)
    path_group[:9] = np.arange(9)
    source_count = rng.integers(1, 40, size=48).astype(np.uint32)
    expect, got = run_both(ctx, ref, eff, mapq, reads, min_noise=1e-4, path_group=path_group, source_count=source_count)
    assert_same(expect, got)
    assert got[4].max() < 9


def test_several_clusters_in_one_call(ctx, ref):
    rng = np.random.default_rng(11)
    parts, read_offset, path_offset = [], [0], [0]
    args = dict(eff=[], mapq=[], align_ptr=[0], score=[], alen=[], flen=[], id_ptr=[0], ids=[])
    expected = []
    for c in range(5):
        eff, mapq, reads = random_reads(rng, n_paths=int(rng.integers(3, 40)), n_reads=int(rng.integers(1, 200)))
        align_ptr, score, alen, flen, id_ptr, ids = [0], [], [], [], [0], []
        for read in reads:
            for (s, l, f, p) in read:
                score.append(s); alen.append(l); flen.append(f); ids.extend(p); id_ptr.append(len(ids))
            align_ptr.append(len(score))
        expected.append(loader.ref_add_path_probs(ref, eff, mapq, align_ptr, score, alen, flen, id_ptr, ids, min_noise_prob=1e-4, **FRAG))
        a0, i0 = len(args["score"]), len(args["ids"])
        args["eff"].extend(eff); args["mapq"].extend(mapq); args["score"].extend(score); args["alen"].extend(alen); args["flen"].extend(flen)
        args["ids"].extend(ids)
        args["align_ptr"].extend(a0 + x for x in align_ptr[1:])
        args["id_ptr"].extend(i0 + x for x in id_ptr[1:])
        read_offset.append(read_offset[-1] + len(reads))
        path_offset.append(path_offset[-1] + len(eff))
    noise, gptr, prob, iptr, ids = ctx.add_path_probs(read_offset, path_offset, args["eff"], args["mapq"], args["align_ptr"], args["score"], args["alen"],
                                                      frag_log_probs(ref, args["flen"]), args["id_ptr"], args["ids"], min_noise_prob=1e-4)
    for c in range(5):
        r0, r1 = read_offset[c], read_offset[c + 1]
        g0, g1 = int(gptr[r0]), int(gptr[r1])
        got = (noise[r0:r1], gptr[r0:r1 + 1] - gptr[r0], prob[g0:g1], iptr[g0:g1 + 1] - iptr[g0], ids[int(iptr[g0]):int(iptr[g1])])
        assert_same(expected[c], got)


def test_malformed_input_is_refused(ctx):
    with pytest.raises(capi.RpvgB200Error, match="out of range"):
        ctx.add_path_probs([0, 1], [0, 2], [3.0, 3.0], [10], [0, 2], [3, -100], [5, 0], None, [0, 1, 1], [7])
    with pytest.raises(capi.RpvgB200Error, match="effective length"):
        ctx.add_path_probs([0, 1], [0, 2], [0.0, 0.0], [10], [0, 2], [3, -100], [5, 0], None, [0, 1, 1], [1])
