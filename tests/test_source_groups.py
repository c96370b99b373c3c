"""findPathSourceGroups of the haplotype-transcripts class (rpvg:src/path_abundance_estimator.cpp:493-546): the drop-in
class gathers the haplotype lists by counting and walks a map that only holds the ids; the groups, their order (the
iteration order of the reference's sparsepp map) and their counts must be the reference's own. Host code only: runs
without a GPU, against the reference's class compiled in oracle/_ref."""

import numpy as np
import pytest

from tests.harness import binding


def structured(rng, n_transcripts, paths_per_transcript, n_haplotypes):
    """1000GP-like: every haplotype id sits on exactly one path of every transcript."""
    lists = [[] for _ in range(n_transcripts * paths_per_transcript)]
    for h in rng.permutation(n_haplotypes):
        for t in range(n_transcripts):
            # a few paths carry most haplotypes
            p = min(paths_per_transcript - 1, int(rng.geometric(0.3)) - 1)
            lists[t * paths_per_transcript + p].append(int(h))
    return lists


def scattered(rng, n_paths, n_ids, per_path, id_space):
    pool = rng.choice(id_space, size=n_ids, replace=False)
    return [[int(x) for x in rng.choice(pool, size=min(n_ids, int(rng.integers(0, per_path + 1))), replace=False)] for _ in range(n_paths)]


@pytest.mark.parametrize("case", ["tiny", "structured-small", "structured-1000gp", "scattered", "sparse-ids", "no-ids"])
def test_source_groups_match_the_reference(b200_harness, reference, case):
    rng = np.random.default_rng(sum(map(ord, case)))

    if case == "tiny":
        lists = [[3, 1], [1], [2, 3], []]
    elif case == "structured-small":
        lists = structured(rng, 2, 6, 60)
    elif case == "structured-1000gp":
        lists = structured(rng, 4, 40, 5008)
    elif case == "scattered":
        lists = scattered(rng, 120, 900, 200, 100000)
    elif case == "sparse-ids":
        # ids far apart: the table over the id range is not used
        lists = scattered(rng, 6, 20, 5, 2**32 - 1)
    else:
        lists = [[], [], []]

    ours = binding.harness_source_groups(b200_harness, lists)
    theirs = binding.harness_source_groups(reference, lists)

    assert ours[0] == theirs[0], "groups (and their order) differ from the reference's"
    assert ours[1] == theirs[1], "group counts differ"
    assert sum(theirs[1]) == len({i for x in lists for i in x})
