import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from rpvg_b200 import capi
from tests.test_collapse_gpu import make_problem, to_csr
seed = int(sys.argv[1]); which = int(sys.argv[2])
rng = np.random.default_rng(seed)
problems = [make_problem(rng, int(rows), int(cols), int(distinct)) for rows, cols, distinct in
            [(1, 1, 1), (40, 3, 7), (300, 6, 40), (2000, 4, 150), (5000, 12, 900), (17, 2, 17)]]
sel = problems if which < 0 else [problems[which]]
ctx = capi.Context(0)
try:
    rc, n, rep, cnt = ctx.collapse(*to_csr(sel))
    print("seed", seed, "problem", which, "ok", n)
except Exception as e:
    print("seed", seed, "problem", which, "FAIL", str(e)[:80])
