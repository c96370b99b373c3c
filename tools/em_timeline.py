"""Per-CTA timeline of one EM pass over the config-2 arena (needs tools/bin/librpvg_b200_timeline.so, built with
`tools/build_variant.sh timeline -DRPVG_B200_EM_TIMELINE`). Prints how busy the SMs were and what ran last.

  python tools/em_timeline.py            # parent: runs the child, parses its device printf lines
"""
import os
import subprocess
import sys
from collections import defaultdict

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    sys.path.insert(0, REPO)
    import ctypes as C
    import numpy as np
    import torch
    from rpvg_b200 import capi, synth

    arena = synth.transcripts_arena()
    stream = torch.cuda.Stream()
    ctx = capi.Context(0, stream.cuda_stream)
    batch = ctx.upload(arena)
    lib = capi.lib()
    rec = np.zeros((65536, 12), dtype=np.uint64)
    n = C.c_uint(0)
    batch.em()
    lib.rpvg_b200_timeline_read(rec.ctypes.data_as(C.c_void_p), 65536, C.byref(n))
    batch.em()
    lib.rpvg_b200_timeline_read(rec.ctypes.data_as(C.c_void_p), 65536, C.byref(n))
    print("=== RUN 2", flush=True)
    for r in rec[: n.value]:
        print("[timeline]", " ".join(str(int(x)) for x in r))
    print("=== END", flush=True)


def parent():
    env = dict(os.environ)
    env["RPVG_B200_LIB"] = os.path.join(REPO, "tools", "bin", "librpvg_b200_timeline.so")
    env["RPVG_TIMELINE_CHILD"] = "1"
    out = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=900)
    text = out.stdout
    if "=== RUN 2" not in text:
        print(text[-2000:], out.stderr[-2000:])
        raise SystemExit("child failed")
    body = text.split("=== RUN 2", 1)[1]
    rows = []
    for line in body.splitlines():
        if line.startswith("[timeline]"):
            f = line.split()[1:]
            rows.append(tuple(int(x) for x in f))
    # threads S rank cluster smid M N t0 t1 t2 its all_shared
    t_min = min(r[7] for r in rows)
    t_max = max(r[9] for r in rows)
    span = (t_max - t_min) / 1e6
    print(f"CTAs {len(rows)}  span {span:.2f} ms")
    busy = defaultdict(float)
    for r in rows:
        busy[r[4]] += (r[9] - r[7]) / 1e6
    sm_ids = sorted(busy)
    tot = sum(busy.values())
    print(f"SMs seen {len(sm_ids)}  sum of CTA time {tot:.1f} ms = {tot / span:.1f} SM-equivalents busy (CTAs can share an SM)")
    # occupancy over time: number of distinct SMs with at least one CTA resident, sampled at 100 points
    samples = 40
    line = []
    for k in range(samples):
        t = t_min + (t_max - t_min) * (k + 0.5) / samples
        sms = {r[4] for r in rows if r[7] <= t < r[9]}
        line.append(len(sms))
    print("SMs with a resident CTA over time:", " ".join(str(x) for x in line))
    # by bin (threads, S)
    bins = defaultdict(list)
    for r in rows:
        bins[(r[0], r[1])].append(r)
    print("bin (threads,S): CTAs, first start, last end (ms from step start), sum CTA ms, prologue share")
    for key in sorted(bins, key=lambda k: (-k[1], -k[0])):
        rs = bins[key]
        s0 = (min(r[7] for r in rs) - t_min) / 1e6
        s1 = (max(r[9] for r in rs) - t_min) / 1e6
        tt = sum(r[9] - r[7] for r in rs) / 1e6
        pro = sum(r[8] - r[7] for r in rs) / 1e6
        print(f"  {key}: {len(rs):6d}  {s0:7.2f} {s1:7.2f}  {tt:9.1f}  {pro / tt:5.2f}")
    # the longest CTAs
    longest = sorted(rows, key=lambda r: r[9] - r[7], reverse=True)[:12]
    print("longest CTAs: threads S rank cluster M N its ms start end us/it shared")
    seen = set()
    for r in sorted(rows, key=lambda r: r[9] - r[7], reverse=True):
        if r[3] in seen:
            continue
        seen.add(r[3])
        d = (r[9] - r[7]) / 1e6
        print(f"  {r[0]} {r[1]} {r[2]} c{r[3]} M{r[5]} N{r[6]} its {r[10]}  {d:.2f} ms  [{(r[7] - t_min) / 1e6:.2f}, {(r[9] - t_min) / 1e6:.2f}]  {1e3 * (r[9] - r[8]) / 1e6 / max(r[10], 1):.2f} us/it  sh{r[11]}")
        if len(seen) >= 25:
            break
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as fh:
            for r in rows:
                fh.write(" ".join(str(x) for x in r) + "\n")


if __name__ == "__main__":
    if os.environ.get("RPVG_TIMELINE_CHILD"):
        child()
    else:
        parent()
