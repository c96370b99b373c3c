"""Offline estimate of the shared-memory wavefronts of the E-step gather (a[col]) for one config-2 cluster, as the kernel
lays it out (rows by first path, stable by length, 32-row chunks), and what a greedy per-step choice of each row's next
entry would give. Usage: python tools/bank_sim.py [target_weight]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rpvg_b200 import synth

def wavefronts(addr):
    """addr: 32 ints (or -1 for inactive lanes): 8-byte word addresses. Two half-warps, conflict degree each."""
    total = 0
    for h in range(2):
        a = addr[16 * h:16 * h + 16]
        a = a[a >= 0]
        if a.size == 0:
            continue
        distinct = np.unique(a)
        banks = distinct % 16
        total += np.bincount(banks, minlength=16).max()
    return total

def main():
    target = int(sys.argv[1]) if len(sys.argv) > 1 else 12000
    arena = synth.transcripts_arena()
    shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
    weight = shapes[:, 2] + shapes[:, 0]
    c = int(np.argmin(np.abs(weight - target)))
    sub = arena.select([c])
    M = sub.n_reads
    rows = [sub.col[int(sub.row_ptr[i]):int(sub.row_ptr[i + 1])].astype(np.int64) for i in range(M)]
    order = sorted(range(M), key=lambda i: (rows[i][0] if len(rows[i]) else 1 << 30))
    rows = [rows[i] for i in order]
    lens = np.array([len(r) for r in rows])
    order2 = np.argsort(-np.minimum(lens, 10 ** 6), kind="stable")
    rows = [rows[i] for i in order2]
    depth = max(8, min(64, 2 * int(lens.sum()) // (32 * 16)))
    print("cluster", c, "M", M, "N", sub.n_paths, "entries", int(lens.sum()), "depth", depth)
    base = greedy = ideal = steps = 0
    for k in range(0, M, 32):
        chunk = [r for r in rows[k:k + 32]]
        if len(chunk[0]) > depth:
            continue   # multi-lane rows: not simulated
        width = len(chunk[0])
        # as laid out: entry t of each row
        for t in range(width):
            addr = np.array([r[t] if t < len(r) else -1 for r in chunk] + [-1] * (32 - len(chunk)))
            base += wavefronts(addr)
            ideal += 2 if (addr >= 0).sum() > 16 else 1
            steps += 1
        # greedy: per step, lanes in order pick an unused entry whose bank is free (or whose address is already taken)
        left = [list(r) for r in chunk]
        for t in range(width):
            addr = -np.ones(32, dtype=np.int64)
            for h in range(2):
                taken_bank = {}
                for lane in range(16 * h, min(16 * h + 16, len(chunk))):
                    cand = left[lane]
                    if not cand:
                        continue
                    pick = None
                    for x in cand:
                        b = x % 16
                        if b not in taken_bank or taken_bank[b] == x:
                            pick = x
                            break
                    if pick is None:
                        pick = cand[0]
                    cand.remove(pick)
                    taken_bank.setdefault(pick % 16, pick)
                    addr[lane] = pick
            greedy += wavefronts(addr)
    print(f"steps {steps}: wavefronts as laid out {base / steps:.2f}, greedy {greedy / steps:.2f}, ideal {ideal / steps:.2f}")

main()


def sigma_sim(target, sigmas=(0, 64, 128, 256, 512, 1024)):
    arena = synth.transcripts_arena()
    shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
    weight = shapes[:, 2] + shapes[:, 0]
    c = int(np.argmin(np.abs(weight - target)))
    sub = arena.select([c])
    M = sub.n_reads
    rows = [sub.col[int(sub.row_ptr[i]):int(sub.row_ptr[i + 1])].astype(np.int64) for i in range(M)]
    order = sorted(range(M), key=lambda i: (rows[i][0] if len(rows[i]) else 1 << 30))
    rows = [rows[i] for i in order]
    entries = sum(len(r) for r in rows)
    for sigma in sigmas:
        if sigma == 0:
            srt = sorted(rows, key=lambda r: -len(r))
        else:
            srt = []
            for w0 in range(0, M, sigma):
                srt += sorted(rows[w0:w0 + sigma], key=lambda r: -len(r))
        steps = gather = 0
        for k in range(0, M, 32):
            chunk = srt[k:k + 32]
            width = max(len(r) for r in chunk)
            for t in range(width):
                addr = np.array([r[t] if t < len(r) else -1 for r in chunk] + [-1] * (32 - len(chunk)))
                gather += wavefronts(addr)
                steps += 1
        total = 3 * steps + gather
        print(f"sigma {sigma:5d}: steps {steps} (padding x{steps * 32 / entries:.2f}) gather {gather / steps:.2f}/step  total wavefronts {total} = {total / entries * 32:.2f} per 32 entries")


if len(sys.argv) > 2:
    sigma_sim(int(sys.argv[1]))


def lex_sim(target):
    arena = synth.transcripts_arena()
    shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
    weight = shapes[:, 2] + shapes[:, 0]
    c = int(np.argmin(np.abs(weight - target)))
    sub = arena.select([c])
    M = sub.n_reads
    rows = [tuple(sub.col[int(sub.row_ptr[i]):int(sub.row_ptr[i + 1])].astype(np.int64)) for i in range(M)]
    print("distinct column sets", len(set(rows)), "of", M, "rows")
    for name, key in (("first col, then length", None), ("length, then whole column list", lambda r: (-len(r), r)),
                      ("length, then last col", lambda r: (-len(r), r[-1] if r else 0, r))):
        if key is None:
            srt = sorted(rows, key=lambda r: (r[0] if r else 1 << 30))
            srt = sorted(srt, key=lambda r: -len(r))
        else:
            srt = sorted(rows, key=key)
        steps = gather = 0
        for k in range(0, M, 32):
            chunk = srt[k:k + 32]
            width = max(len(r) for r in chunk)
            for t in range(width):
                addr = np.array([r[t] if t < len(r) else -1 for r in chunk] + [-1] * (32 - len(chunk)))
                gather += wavefronts(addr)
                steps += 1
        print(f"{name}: steps {steps} gather {gather / steps:.2f}/step")


if len(sys.argv) > 2 and sys.argv[2] == "lex":
    lex_sim(int(sys.argv[1]))
