"""Runs the EM of one config-2 cluster (closest to the given weight) once: a short target for ncu."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rpvg_b200 import capi, synth  # noqa: E402

target = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
group = int(sys.argv[2]) if len(sys.argv) > 2 else 1
arena = synth.transcripts_arena()
shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
weight = shapes[:, 2] + shapes[:, 0]
c = int(np.argmin(np.abs(weight - target)))
sub = arena.select([c])
os.environ["RPVG_B200_EM_SLICE_WEIGHT"] = str(max(64, int(weight[c] // group) - 1 if group > 1 else 10 ** 9))
os.environ["RPVG_B200_EM_MAX_GROUP"] = str(group)
ctx = capi.Context(0)
ab, nz, tc, its = ctx.em_abundances(sub)
print("cluster", c, "weight", weight[c], "N", shapes[c][1], "its", its[0], "sum", ab.sum())
