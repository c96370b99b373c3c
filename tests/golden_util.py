"""Loading of tests/golden/*.npz (frozen output of the reference's own code, see tests/golden/make_golden.py)."""
import os

import numpy as np

from rpvg_b200.arena import ClusterArena

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLDEN, name))
    fields = {k[len("arena_"):]: z[k] for k in z.files if k.startswith("arena_")}
    return ClusterArena(**fields), z


def cluster_results(z, prefix, c):
    """(sets, posteriors, abundances, noise_count, total_count) of cluster c under the given result prefix."""
    s0, s1 = int(z[prefix + "set_ptr"][c]), int(z[prefix + "set_ptr"][c + 1])
    sizes = z[prefix + "set_size"]
    start = int(sizes[:s0].sum())
    sets = []
    for k in range(s0, s1):
        sets.append(tuple(int(x) for x in z[prefix + "set_ids"][start:start + int(sizes[k])]))
        start += int(sizes[k])
    p0, p1 = int(z[prefix + "post_ptr"][c]), int(z[prefix + "post_ptr"][c + 1])
    a0, a1 = int(z[prefix + "abund_ptr"][c]), int(z[prefix + "abund_ptr"][c + 1])
    return sets, z[prefix + "posteriors"][p0:p1], z[prefix + "abundances"][a0:a1], float(z[prefix + "noise_count"][c]), float(z[prefix + "total_count"][c])
