"""Offline model of the gather wavefronts of both EM phases for one single-CTA config-2 cluster under the kernel's
sliced-ELL layout, as laid out and with the lane-private greedy reordering (each lane permutes its own entries so that
the 16 lanes of a half-warp hit different banks at each step). python tools/bank_sim2.py [weight]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rpvg_b200 import synth


def bucket_upper(n):
    if n <= 32:
        return n
    return 1 << int(np.ceil(np.log2(n)))


def lanes_log2(upper, depth):
    want = -(-upper // depth)
    l2 = 0
    while l2 < 5 and (1 << l2) < want:
        l2 += 1
    return l2


def wavefronts(addr):
    total = 0
    for h in range(2):
        a = addr[16 * h:16 * h + 16]
        a = a[a >= 0]
        if a.size:
            total += np.bincount(np.unique(a) % 16, minlength=16).max()
    return total


def layout(items, depth):
    """items: list of index arrays, sorted by length descending. Yields chunks: 32 per-lane entry lists + width."""
    p = 0
    n = len(items)
    while p < n:
        l2 = lanes_log2(bucket_upper(len(items[p])), depth)
        L = 1 << l2
        count = min(32 >> l2, n - p)
        lanes = [[] for _ in range(32)]
        width = 0
        for r in range(count):
            ent = items[p + r]
            for sub in range(L):
                lanes[r * L + sub] = list(ent[sub::L])
            width = max(width, -(-len(ent) // L))
        yield lanes, width
        p += count


def cost(chunks, greedy):
    steps = total = 0
    for lanes, width in chunks:
        left = [list(l) for l in lanes]
        for t in range(width):
            addr = -np.ones(32, dtype=np.int64)
            if not greedy:
                for i in range(32):
                    if t < len(lanes[i]):
                        addr[i] = lanes[i][t]
            else:
                for h in range(2):
                    taken = set()
                    for i in range(16 * h, 16 * h + 16):
                        if not left[i]:
                            continue
                        pick = next((x for x in left[i] if x % 16 not in taken), left[i][0])
                        left[i].remove(pick)
                        taken.add(pick % 16)
                        addr[i] = pick
            total += wavefronts(addr)
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
    order = sorted(range(M), key=lambda i: (rows[i][0] if len(rows[i]) else 1 << 30))   # build kernel: by first path
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
    swz = lambda a: a ^ ((a >> 4) & 15)
    col_items_swz = [swz(a) for a in col_items]
    row_items_swz = [swz(a) for a in row_items]
    for name, items in (("E-step gather a[col]", row_items), ("E-step, swizzled a", row_items_swz), ("M-step gather w[row]", col_items), ("M-step, swizzled w", col_items_swz)):
        ch = list(layout(items, depth))
        s0, w0 = cost(ch, False)
        s1, w1 = cost(ch, True)
        print(f"  {name}: {s0} steps, wavefronts per step as laid out {w0 / s0:.2f}, lane-private greedy {w1 / s1:.2f} (ideal ~2)")


main()
