#!/usr/bin/env python
"""Headline benchmark: EM throughput (nnz x iterations / s) of the transcripts-model hot path on config 2 of
BASELINE.json ("transcripts model EM on synthetic clusters: 1M read classes, 200k paths, power-law cluster sizes").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One process per GPU (torchrun for N > 1). A step = one pass of the hot path (matrix build + EM of every cluster) over
one synthetic arena of the named shape. Clusters are independent, so ranks are sharded with no data-path collective:
every rank owns an arena of the full config-2 shape (seed + rank) -> "scaling": "weak"; value = units of all ranks /
max-over-ranks device time.

* value  : inputs resident in HBM when the timed region starts (CUDA events on the launching stream around each step;
           L2 flushed between steps because the 72 MB arena is smaller than the 126 MB L2).
* e2e    : the same metric through the host-facing C-ABI call (pinned host buffers in, H2D + build + EM + D2H inside
           the timed region).
* roofline : the EM kernel's algorithmic bytes (12 nnz_path + 20 M + 16 (N+1) per cluster per iteration)
           / its own device time, against MEASURED_PEAKS.json's HBM GB/s.
* cpu_baseline / --impl reference : the reference's own estimator objects (oracle/_ref, rpvg's unmodified translation
           units, OpenMP over clusters like rpvg:src/main.cpp:829) on a bounded sample of the same arena.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "em_nnz_iterations_per_second"
UNIT = "nnz*it/s"
WORKLOAD = "config2: transcripts-model EM, 1M read classes, 200k paths, power-law cluster sizes 1-5000 (synthetic, seeds 20240603..)"


def measured_peak_hbm():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._thread = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=6)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
                for name, v in zip(names, s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def reference_sample(arena, max_dense_entries=1.5e6, max_clusters=400, max_total_dense=1.2e8):
    """A bounded sample of the arena's clusters for the CPU arm. The reference is dense: a cluster
    costs M x (N+1) x iterations, so the sample keeps clusters whose dense matrix has at most ``max_dense_entries``
    entries, evenly spaced over the size-sorted list, within a total dense budget."""
    shapes = [arena.cluster_shape(c) for c in range(arena.n_clusters)]
    dense = np.array([m * (n + 1) for m, n, _ in shapes], dtype=np.float64)
    eligible = np.nonzero((dense <= max_dense_entries) & (np.array([m for m, _, _ in shapes]) > 0))[0]
    eligible = eligible[np.argsort(-dense[eligible], kind="stable")]
    picks = eligible[np.linspace(0, len(eligible) - 1, min(max_clusters, len(eligible))).astype(int)]
    picks = np.unique(picks)
    picks = picks[np.argsort(-dense[picks], kind="stable")]
    keep, total = [], 0.0
    for c in picks[::-1]:   # smallest first until the budget is used, so the sample is never empty
        if total + dense[c] > max_total_dense and keep:
            continue
        keep.append(int(c))
        total += dense[c]
    return sorted(keep)


def host_threads() -> int:
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1: the CPU arm must not inherit that)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference_sample(arena, sample_ids, threads=0):
    """Times the reference's own PathAbundanceEstimator (oracle/_ref) — or, if that library is absent, the CPU port — on
    the sample. Returns (nnz*it/s, seconds, kind, cores)."""
    from oracle import loader
    threads = threads or host_threads()
    sub = arena.select(sample_ids)
    _, _, _, its, _ = loader.port_em_abundances(sub, num_threads=threads)     # iteration counts (the reference does not report them)
    work = 0.0
    for k in range(sub.n_clusters):
        m, _, nnz_path = sub.cluster_shape(k)
        work += float(its[k]) * (nnz_path + m)
    if loader.reference_available():
        lib = loader.reference()
        cfg = loader.make_config(loader.MODEL_TRANSCRIPTS)
        seconds, _ = loader.harness_estimate_batch(lib, cfg, sub, rng_seed=1, num_threads=threads)
        return work / seconds, seconds, "reference", threads
    _, _, _, _, seconds = loader.port_em_abundances(sub, num_threads=threads)
    return work / seconds, seconds, "port", threads


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; a scaled run is not a bench number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    from rpvg_b200 import synth

    n_paths = max(200, int(200_000 * args.scale))
    n_reads = max(1000, int(1_000_000 * args.scale))
    n_max = max(20, int(5000 * min(1.0, args.scale * 4))) if args.scale < 1 else 5000
    config = {"workload": WORKLOAD if args.scale == 1.0 else f"DEBUG scale={args.scale} of config2",
              "read_classes_per_gpu": n_reads, "paths_per_gpu": n_paths, "max_em_its": 10000, "max_rel_em_conv": 0.001,
              "sharding": f"{world} arena(s) of this shape; clusters dealt to {world} rank(s) by LPT on nnz x E[iterations], no collective", "l2": "flushed between timed steps (256 MiB write)"}

    # ------------------------------------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        arena = synth.transcripts_arena(seed=synth.BASE_SEED + 2, n_paths=n_paths, n_reads=n_reads, n_max=n_max)
        sample = reference_sample(arena)
        for _ in range(args.warmup):
            run_reference_sample(arena, sample[: max(1, len(sample) // 8)])
        values, secs = [], []
        kind, cores = "reference", 0
        for _ in range(args.steps):
            v, s, kind, cores = run_reference_sample(arena, sample)
            values.append(v)
            secs.append(s)
        value = float(np.sum([v * s for v, s in zip(values, secs)]) / np.sum(secs))
        sample_desc = f"{len(sample)} of {arena.n_clusters} clusters (dense M x (N+1) <= 1.5e6 each), all host threads"
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * float(np.mean(secs)), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": config, "impl": "reference",
                "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_desc},
                "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------------------------------------------ B200 arm
    import torch
    import torch.distributed as dist
    from rpvg_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if world == 1:
        arena = synth.transcripts_arena(seed=synth.BASE_SEED + 2, n_paths=n_paths, n_reads=n_reads, n_max=n_max)
    else:
        # weak scaling: the job is `world` arenas of the config-2 shape; its clusters are dealt to the ranks by LPT on the
        # nnz x E[iterations] cost model (rpvg_b200/multi_gpu.py). No data-path collective: clusters are independent.
        from rpvg_b200 import multi_gpu
        from rpvg_b200.arena import concat
        job = concat([synth.transcripts_arena(seed=synth.BASE_SEED + 2 + r, n_paths=n_paths, n_reads=n_reads, n_max=n_max) for r in range(world)])
        arena = job.select(multi_gpu.lpt_partition(multi_gpu.cluster_costs(job), world)[rank])
        del job

    # pinned host copies of the arena: the e2e leg uploads from these every step
    class Pinned:
        pass

    pinned = Pinned()
    for name in ("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count"):
        a = getattr(arena, name)
        signed = {np.dtype(np.uint64): np.int64, np.dtype(np.uint32): np.int32}.get(a.dtype)
        t = torch.from_numpy(a.view(signed) if signed else a).pin_memory()
        setattr(pinned, name, t)
    h2d_bytes = arena.input_bytes()

    out_ab = torch.empty(arena.n_paths, dtype=torch.float64).pin_memory()
    out_noise = torch.empty(arena.n_clusters, dtype=torch.float64).pin_memory()
    out_total = torch.empty(arena.n_clusters, dtype=torch.float64).pin_memory()
    out_its = torch.empty(arena.n_clusters, dtype=torch.int32).pin_memory()
    d2h_bytes = out_ab.numel() * 8 + out_noise.numel() * 8 + out_total.numel() * 8 + out_its.numel() * 4

    # a dedicated (non-default) stream shared by torch (events, L2 flush) and the library (kernels, copies)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx = capi.Context(local_rank, stream.cuda_stream)
    params = capi.default_em_params()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg ("value")
    batch = ctx.upload(pinned)
    for _ in range(args.warmup):
        batch.em(params)
    batch.download(out_ab, out_noise, out_total, out_its)
    work, alg_bytes = batch.em_work()

    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    launches_before = ctx.launches
    step_ms, em_ms, build_ms = [], [], []
    for _ in range(args.steps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        batch.em(params)
        e1.record(stream)
        e1.synchronize()
        step_ms.append(e0.elapsed_time(e1))
        b_ms, k_ms = batch.em_timings()
        build_ms.append(b_ms)
        em_ms.append(k_ms)
    launches = ctx.launches - launches_before
    barrier()
    clocks = sampler.stop()
    batch.download(out_ab, out_noise, out_total, out_its)
    checksum = float(out_ab.sum())
    batch.free()

    # ---- end-to-end leg ("e2e"): host buffers -> C ABI -> host buffers
    for _ in range(min(2, args.warmup)):
        b = ctx.upload(pinned)
        b.em(params)
        b.download(out_ab, out_noise, out_total, out_its)
        b.free()
    barrier()
    e2e_ms = []
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        b = ctx.upload(pinned)
        b.em(params)
        b.download(out_ab, out_noise, out_total, out_its)
        e2e_ms.append(1e3 * (time.perf_counter() - t0))
        b.free()
    barrier()

    total_ms = float(np.sum(step_ms))
    total_e2e_ms = float(np.sum(e2e_ms))
    stats = torch.tensor([total_ms, total_e2e_ms], dtype=torch.float64, device="cuda")
    works = torch.tensor([work], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(works, op=dist.ReduceOp.SUM)
    max_ms, max_e2e_ms = [float(x) for x in stats.cpu()]
    all_work = float(works.cpu()[0])

    if rank == 0:
        value = all_work * args.steps / (max_ms * 1e-3)
        e2e_value = all_work * args.steps / (max_e2e_ms * 1e-3)
        peak, peak_src = measured_peak_hbm()
        em_ms_mean = float(np.mean(em_ms))
        achieved = alg_bytes / (em_ms_mean * 1e-3) / 1e9
        # DRAM bytes the EM launches of one step really move (ncu, same workload): the slices live in shared memory, so
        # this is a small fraction of the algorithmic bytes the HBM roofline is quoted on (profiles/r01d_*).
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01d_em_traffic.json")) as fh:
                measured = json.load(fh)
            if world == 1:
                traffic, traffic_src = measured["em_dram_bytes_per_step"], measured["source"]
        except OSError:
            pass
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                        "ms_per_step": max_e2e_ms / args.steps},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "emSellKernel (all bins, concurrent streams)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                             "limiter": "shared-memory data pipe: 76% busy over a whole single-cluster launch (ncu, profiles/r01c_ncu_single_cluster_metrics.txt); the rest is barriers",
                             "shared_pipe_busy_frac": 0.76,
                             "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": em_ms_mean, "build_ms": float(np.mean(build_ms)),
                             "kernel_share_of_step": em_ms_mean / (max_ms / args.steps)},
                "inference_stage_wall_ms": max_e2e_ms / args.steps,
                "work_per_step": {"nnz_iterations": all_work, "nnz": arena.nnz, "clusters": arena.n_clusters,
                                  "max_iterations": int(out_its.max()), "checksum": checksum}}
        if not args.no_cpu_baseline:
            sample = reference_sample(arena, max_clusters=200, max_total_dense=4e7)
            v, s, kind, cores = run_reference_sample(arena, sample)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "seconds": s,
                                    "sample": f"{len(sample)} of {arena.n_clusters} clusters (dense M x (N+1) <= 1.5e6 each)"}
        print(json.dumps(line))

    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
