"""Offline model of the gather wavefronts with K copies of the gathered vector at different bank offsets (each entry may
read any copy) on top of the lane-private reordering. python tools/bank_sim3.py [weight]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rpvg_b200 import synth
from tools.bank_sim2 import bucket_upper, lanes_log2, layout  # noqa


def cost(chunks, copies, use_bcast):
    steps = total = 0
    shift = 16 // copies
    for lanes, width in chunks:
        left = [list(l) for l in lanes]
        for t in range(width):
            tot_step = 0
            for h in range(2):
                taken = {}   # bank -> address
                load = np.zeros(16, dtype=np.int64)
                for i in range(16 * h, 16 * h + 16):
                    if not left[i]:
                        continue
                    best = None
                    for x in left[i]:
                        for c in range(copies):
                            b = (x + c * shift) % 16
                            addr = (x, c)
                            if b not in taken:
                                best = (x, c, b); break
                            if use_bcast and addr in taken[b]:
                                best = (x, c, b); break
                        if best: break
                    if best is None:
                        # least loaded bank among options of first entry
                        x = left[i][0]
                        opts = [((x + c * shift) % 16, c) for c in range(copies)]
                        b, c = min(opts, key=lambda o: len(taken[o[0]]))
                        best = (x, c, b)
                    x, c, b = best
                    left[i].remove(x)
                    taken.setdefault(b, set()).add((x, c))
                tot_step += max((len(v) for v in taken.values()), default=0)
            total += tot_step
            steps += 1
    return steps, total


def main():
    target = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
    arena = synth.transcripts_arena()
    shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
    weight = shapes[:, 2] + shapes[:, 0]
    c = int(np.argmin(np.abs(weight - target)))
    sub = arena.select([c])
    M, N = sub.n_reads, sub.n_paths
    rows = [sub.col[int(sub.row_ptr[i]):int(sub.row_ptr[i + 1])].astype(np.int64) for i in range(M)]
    order = sorted(range(M), key=lambda i: (rows[i][0] if len(rows[i]) else 1 << 30))
    rows = [rows[i] for i in order]
    z = sum(len(r) for r in rows)
    depth = max(8, min(64, 2 * z // (32 * 16)))
    row_order = sorted(range(M), key=lambda i: -bucket_upper(len(rows[i])))
    row_pos = np.empty(M, dtype=np.int64)
    row_pos[row_order] = np.arange(M)
    row_items = [rows[i] for i in row_order]
    cols = [[] for _ in range(N)]
    for i in range(M):
        for j in rows[i]:
            cols[j].append(row_pos[i])
    col_items = sorted([np.array(cj, dtype=np.int64) for cj in cols if cj], key=lambda a: -bucket_upper(len(a)))
    print(f"cluster {c}: M {M} N {N} entries {z} depth {depth}")
    for name, items in (("E-step gather a[col]", row_items), ("M-step gather w[row]", col_items)):
        ch = list(layout(items, depth))
        for copies in (1, 2, 4):
            for bc in (False, True):
                s, w = cost(ch, copies, bc)
                print(f"  {name}: copies {copies} broadcast-aware {bc}: {w / s:.2f} wavefronts per step")


main()
