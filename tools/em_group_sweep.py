"""Per-iteration EM time of single clusters as a function of the thread-block-cluster size (tuning aid, GPU only)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rpvg_b200 import capi, synth  # noqa: E402

arena = synth.transcripts_arena()
shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
weight = shapes[:, 2] + shapes[:, 0]
ctx = capi.Context(0)

targets = [190000, 118000, 84000, 48000, 26000, 13600, 9000, 5000, 2000, 500, 60]
for t in targets:
    c = int(np.argmin(np.abs(weight - t)))
    sub = arena.select([c])
    line = f"cluster {c} weight {weight[c]} N {shapes[c][1]} M {shapes[c][0]}:"
    for group in (16, 8, 4, 2, 1):
        if weight[c] < 1500 * group and group > 1:
            continue
        os.environ["RPVG_B200_EM_SLICE_WEIGHT"] = str(max(64, int(weight[c] // group) - 1 if group > 1 else 10 ** 9))
        os.environ["RPVG_B200_EM_MAX_GROUP"] = str(group)
        batch = ctx.upload(sub)
        batch.em()
        batch.em()
        ab = np.empty(sub.n_paths); nz = np.empty(1); tc = np.empty(1); its = np.empty(1, dtype=np.uint32); its1 = np.empty(1, dtype=np.uint32)
        batch.download(ab, nz, tc, its)
        _, em_ms = batch.em_timings()
        batch.em(capi.default_em_params(max_em_its=1))
        batch.download(ab, nz, tc, its1)
        _, pro_ms = batch.em_timings()
        batch.free()
        line += f"  S={group}: {em_ms:.2f} ms (prologue {pro_ms:.2f}) / {int(its[0])} its = {1e3 * (em_ms - pro_ms) / max(1, int(its[0])):.2f} us"
    print(line, flush=True)
