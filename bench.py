#!/usr/bin/env python
"""Headline benchmark: EM throughput (nnz x iterations / s) of the transcripts-model hot path on config 2 of
BASELINE.json ("transcripts model EM on synthetic clusters: 1M read classes, 200k paths, power-law cluster sizes").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c1|c2|c3|c4|c5]

--config (default c2, the headline) names the BASELINE.json configuration that runs; c1/c3/c4/c5 run the model classes
through the batched driver and report the inference-stage wall time (bench_models.py).

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
           units, OpenMP over clusters like rpvg:src/main.cpp:829) on a bounded sample of the same arena, sized from the
           clusters' own cost (dense entries x iterations) so that every host thread stays busy; `same_work` is the GPU on
           exactly that sample; `cpu_port` is the sparse CPU restatement (oracle/rpvg_oracle.cpp) on the WHOLE arena, so the
           share of the ratio that is sparse-vs-dense algorithmics is visible next to the hardware's.
* strong   : (N > 1) one config-2 arena, and a job of 8 of them, dealt over the N ranks by LPT: fixed total work.
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


def reference_sample(arena, iterations, threads, target_seconds=12.0, dense_rate=2.5e8, seed=7):
    """A bounded sample of the arena's clusters for the CPU arm. The reference is dense: a cluster costs
    M x (N + 1) x iterations there. The sample is drawn at random (seeded) among the clusters whose cost is at most 1/(4 x
    threads) of the budget — so that rpvg's dynamic largest-first loop keeps every thread busy to the end — until about
    ``target_seconds`` x threads of single-thread work (at ``dense_rate`` dense entries x iterations per second) is reached."""
    shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)], dtype=np.float64)
    cost = shapes[:, 0] * (shapes[:, 1] + 1) * np.asarray(iterations, dtype=np.float64)
    budget = target_seconds * threads * dense_rate
    eligible = np.nonzero((cost > 0) & (cost <= budget / (4.0 * threads)))[0]
    order = np.random.default_rng(seed).permutation(eligible)
    keep, total = [], 0.0
    for c in order:
        if total + cost[c] > budget:
            continue
        keep.append(int(c))
        total += cost[c]
    return sorted(keep), float(total)


def host_threads() -> int:
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1: the CPU arm must not inherit that)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_seconds():
    t = os.times()
    return t.user + t.system


def run_reference_sample(sub, its, threads=0):
    """Times the reference's own PathAbundanceEstimator (oracle/_ref) — or, if that library is absent, the CPU port — on
    the clusters of ``sub`` (iteration counts ``its`` from the port: the reference does not report them).
    Returns (nnz*it/s, seconds, kind, cores, host-core utilisation during the run)."""
    from oracle import loader
    threads = threads or host_threads()
    work = 0.0
    for k in range(sub.n_clusters):
        m, _, nnz_path = sub.cluster_shape(k)
        work += float(its[k]) * (nnz_path + m)
    c0 = cpu_seconds()
    if loader.reference_available():
        lib = loader.reference()
        cfg = loader.make_config(loader.MODEL_TRANSCRIPTS)
        seconds, _ = loader.harness_estimate_batch(lib, cfg, sub, rng_seed=1, num_threads=threads)
        kind = "reference"
    else:
        _, _, _, _, seconds = loader.port_em_abundances(sub, num_threads=threads)
        kind = "port"
    busy = (cpu_seconds() - c0) / max(seconds * threads, 1e-9)
    return work / seconds, seconds, kind, threads, busy


def run_port_full(arena, threads=0, max_em_its=10000):
    """The sparse CPU restatement on the whole arena, all host threads: (nnz*it/s, seconds, iterations)."""
    from oracle import loader
    threads = threads or host_threads()
    _, _, _, its, seconds = loader.port_em_abundances(arena, max_em_its=max_em_its, num_threads=threads)
    work = 0.0
    for c in range(arena.n_clusters):
        m, _, nnz_path = arena.cluster_shape(c)
        work += float(its[c]) * (nnz_path + m)
    return work / seconds, seconds, its


def pin_arena(torch, arena):
    """Pinned host copies of an arena's arrays (the e2e legs upload from these)."""
    class Pinned:
        pass
    pinned = Pinned()
    for name in ("read_offset", "path_offset", "row_ptr", "col", "prob", "noise", "count"):
        a = getattr(arena, name)
        signed = {np.dtype(np.uint64): np.int64, np.dtype(np.uint32): np.int32}.get(a.dtype)
        setattr(pinned, name, torch.from_numpy(a.view(signed) if signed else a).pin_memory())
    return pinned


def pinned_outputs(torch, arena):
    return (torch.empty(arena.n_paths, dtype=torch.float64).pin_memory(), torch.empty(arena.n_clusters, dtype=torch.float64).pin_memory(),
            torch.empty(arena.n_clusters, dtype=torch.float64).pin_memory(), torch.empty(arena.n_clusters, dtype=torch.int32).pin_memory())


def time_resident(torch, ctx, stream, flush, params, pinned, outs, steps, warmup):
    """EM on an arena resident in HBM: per-step device time (CUDA events on the launching stream, L2 flushed before every
    step). Returns (step_ms list, em_ms list, build_ms list, work, algorithmic bytes, launches)."""
    batch = ctx.upload(pinned)
    for _ in range(warmup):
        batch.em(params)
    batch.download(*outs)
    work, alg_bytes = batch.em_work()
    launches_before = ctx.launches
    step_ms, em_ms, build_ms = [], [], []
    for _ in range(steps):
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
    batch.download(*outs)
    batch.free()
    return step_ms, em_ms, build_ms, work, alg_bytes, launches


def time_e2e(torch, ctx, flush, params, pinned, outs, steps, warmup):
    """Host buffers -> C ABI -> host buffers: upload + build + EM + download per step, wall clock around synchronous calls."""
    for _ in range(warmup):
        b = ctx.upload(pinned)
        b.em(params)
        b.download(*outs)
        b.free()
    times = []
    for _ in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        b = ctx.upload(pinned)
        b.em(params)
        b.download(*outs)
        times.append(1e3 * (time.perf_counter() - t0))
        b.free()
    return times


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c4", "c5", "giant"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; a scaled run is not a bench number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the fixed-work (strong scaling) legs")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.config not in ("c2", "giant"):
        import bench_models
        return bench_models.run(args, rank, local_rank, world, ClockSampler, host_threads)

    from rpvg_b200 import synth

    n_paths = max(200, int(200_000 * args.scale))
    n_reads = max(1000, int(1_000_000 * args.scale))
    n_max = max(20, int(5000 * min(1.0, args.scale * 4))) if args.scale < 1 else 5000
    config = {"workload": WORKLOAD if args.scale == 1.0 else f"DEBUG scale={args.scale} of config2",
              "read_classes_per_gpu": n_reads, "paths_per_gpu": n_paths, "max_em_its": 10000, "max_rel_em_conv": 0.001,
              "sharding": f"{world} arena(s) of this shape; clusters dealt to {world} rank(s) by LPT on nnz x E[iterations], no collective", "l2": "flushed between timed steps (256 MiB write)"}

    giant = args.config == "giant"
    max_em_its = 10000

    if giant:
        # one path cluster larger than the L2 cache in both orientations: the HBM-streaming regime (emStreamKernel)
        giant_paths, giant_reads, max_em_its = max(2000, int(20_000 * args.scale)), max(20000, int(4_000_000 * args.scale)), 200
        config = {"workload": ("giant: ONE path cluster of 20k paths x 4M read classes (~20M entries: 480 MB in the two orientations, beyond the 126 MB L2), "
                               "EM capped at 200 iterations; not a BASELINE.json configuration: the streaming regime north_star asks for") if args.scale == 1.0
                              else f"DEBUG scale={args.scale} of the giant cluster",
                  "read_classes_per_gpu": giant_reads, "paths_per_gpu": giant_paths, "max_em_its": max_em_its, "max_rel_em_conv": 0.001,
                  "sharding": f"{world} replica(s): a single cluster does not shard", "l2": "the cluster is larger than L2; flushed between timed steps all the same"}

    def headline_arena(k=0):
        if giant:
            return synth.transcripts_arena(seed=synth.BASE_SEED + 7 + k, n_reads=giant_reads, sizes=[giant_paths], reads_cap=1e12)
        return synth.transcripts_arena(seed=synth.BASE_SEED + 2 + k, n_paths=n_paths, n_reads=n_reads, n_max=n_max)

    # ------------------------------------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        arena = headline_arena()
        threads = host_threads()
        port_value, port_seconds, its = run_port_full(arena, threads)           # iteration counts + the sparse CPU number
        sample, _ = reference_sample(arena, its, threads)
        sub = arena.select(sample)
        sub_its = its[sample]
        small = list(range(0, sub.n_clusters, 8))
        for _ in range(args.warmup):
            run_reference_sample(sub.select(small), sub_its[small], threads)
        values, secs, busy = [], [], []
        kind, cores = "reference", threads
        for _ in range(args.steps):
            v, s_, kind, cores, b_ = run_reference_sample(sub, sub_its, threads)
            values.append(v)
            secs.append(s_)
            busy.append(b_)
        value = float(np.sum([v * s_ for v, s_ in zip(values, secs)]) / np.sum(secs))
        sample_desc = (f"{len(sample)} of {arena.n_clusters} clusters drawn at random among those costing <= 1/(4 x threads) of the budget "
                       f"(dense M x (N+1) x iterations), all host threads")
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * float(np.mean(secs)), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": config, "impl": "reference",
                "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_desc, "host_core_utilisation": float(np.mean(busy))},
                "cpu_port": {"value": port_value, "unit": UNIT, "cores": threads, "kind": "port", "seconds": port_seconds,
                             "sample": "the whole arena (sparse CPU restatement, oracle/rpvg_oracle.cpp)"},
                "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------------------------------------------ B200 arm
    import torch
    import torch.distributed as dist
    from rpvg_b200 import capi, multi_gpu
    from rpvg_b200.arena import concat

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    first = headline_arena()
    if world == 1:
        arena = first
    else:
        # weak scaling: the job is `world` arenas of the config-2 shape; its clusters are dealt to the ranks by LPT on the
        # nnz x E[iterations] cost model (rpvg_b200/multi_gpu.py). No data-path collective: clusters are independent.
        job = concat([first] + [headline_arena(r) for r in range(1, world)])
        arena = job.select(multi_gpu.lpt_partition(multi_gpu.cluster_costs(job), world)[rank])
        del job

    pinned = pin_arena(torch, arena)
    h2d_bytes = arena.input_bytes()
    outs = pinned_outputs(torch, arena)
    d2h_bytes = outs[0].numel() * 8 + outs[1].numel() * 8 + outs[2].numel() * 8 + outs[3].numel() * 4

    # a dedicated (non-default) stream shared by torch (events, L2 flush) and the library (kernels, copies)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx = capi.Context(local_rank, stream.cuda_stream)
    params = capi.default_em_params(max_em_its=max_em_its)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(values):
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def sum_over_ranks(values):
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t.cpu()]

    # ---- device-resident leg ("value")
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    step_ms, em_ms, build_ms, work, alg_bytes, launches = time_resident(torch, ctx, stream, flush, params, pinned, outs, args.steps, args.warmup)
    barrier()
    clocks = sampler.stop()
    checksum = float(outs[0].sum())
    max_its = int(outs[3].max())
    its_all = outs[3].numpy().astype(np.int64).copy()

    # ---- end-to-end leg ("e2e"): host buffers -> C ABI -> host buffers
    barrier()
    e2e_ms = time_e2e(torch, ctx, flush, params, pinned, outs, args.steps, min(2, args.warmup))
    barrier()

    max_ms, max_e2e_ms = max_over_ranks([float(np.sum(step_ms)), float(np.sum(e2e_ms))])
    all_work = sum_over_ranks([work])[0]

    # ---- fixed total work over the N ranks (strong scaling): one arena, and a job of 8 arenas, dealt by LPT
    strong = None
    if world > 1 and not args.no_strong and args.scale == 1.0 and not giant:
        strong = {}
        for name, n_arenas in (("one_arena", 1), ("eight_arenas", 8)):
            job = first if n_arenas == 1 else concat([first] + [headline_arena(r) for r in range(1, n_arenas)])
            share = job.select(multi_gpu.lpt_partition(multi_gpu.cluster_costs(job), world)[rank])
            sp, so = pin_arena(torch, share), pinned_outputs(torch, share)
            barrier()
            ms, _, _, w_share, _, _ = time_resident(torch, ctx, stream, flush, params, sp, so, 3, 2)
            barrier()
            ms_n = max_over_ranks([float(np.mean(ms))])[0]
            ms_1 = 0.0
            if rank == 0:   # the same job on one GPU, measured in the same run
                jp, jo = pin_arena(torch, job), pinned_outputs(torch, job)
                ms1, _, _, _, _, _ = time_resident(torch, ctx, stream, flush, params, jp, jo, 3, 2)
                ms_1 = float(np.mean(ms1))
                del jp, jo
            barrier()
            ms_1 = max_over_ranks([ms_1])[0]
            strong[name] = {"ms_n1": ms_1, f"ms_n{world}": ms_n, "speedup": ms_1 / ms_n, "work_nnz_iterations": sum_over_ranks([w_share])[0]}
            del job, share, sp, so

    if rank == 0:
        value = all_work * args.steps / (max_ms * 1e-3)
        e2e_value = all_work * args.steps / (max_e2e_ms * 1e-3)
        peak, peak_src = measured_peak_hbm()
        em_ms_mean = float(np.mean(em_ms))
        achieved = alg_bytes / (em_ms_mean * 1e-3) / 1e9
        # DRAM bytes the EM launches of one step really move (ncu, same workload, round-1 build): the slices live in shared
        # memory, so this is a small fraction of the algorithmic bytes the HBM roofline is quoted on.
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "r01d_em_traffic.json")) as fh:
                measured = json.load(fh)
            if world == 1:
                traffic, traffic_src = measured["em_dram_bytes_per_step"], measured["source"] + " (ncu capture of the round-1 build, not of this run)"
        except OSError:
            pass
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                        "ms_per_step": max_e2e_ms / args.steps},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "emSellRunKernel + emSellBuildKernel + emDenseKernel (all bins, concurrent streams)", "achieved": achieved,
                             "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                             "meaning": "algorithmic-bytes equivalent (12 nnz_path + 20 M + 16 (N+1) per cluster and iteration / EM device time): the slices are "
                                        "resident in shared memory, real DRAM traffic is ~0.2% of it; the binding resource is the shared-memory data pipe "
                                        "(profiles/r02c_ncu_single_cluster.md: 74% busy over a whole single-cluster launch, 9.6 wavefronts per 32 entries)",
                             "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": em_ms_mean, "build_ms": float(np.mean(build_ms)),
                             "kernel_share_of_step": em_ms_mean / (max_ms / args.steps)},
                "inference_stage_wall_ms": max_e2e_ms / args.steps,
                "work_per_step": {"nnz_iterations": all_work, "nnz": arena.nnz, "clusters": arena.n_clusters,
                                  "max_iterations": max_its, "checksum": checksum}}
        if strong:
            line["strong"] = strong

        if giant:
            its = float(max_its)
            reads, paths, entries = float(arena.n_reads), float(arena.n_paths), float(arena.nnz_path)
            line["roofline"]["kernel"] = "emStreamKernel (cooperative, whole GPU)"
            line["roofline"]["meaning"] = ("algorithmic bytes (12 nnz_path + 20 M + 16 (N+1) per iteration, the matrix counted once) / EM device time. The kernel reads "
                                           "both orientations of the matrix every iteration (deterministic: no atomics), see bytes_read_per_iteration")
            line["roofline"]["bytes_read_per_iteration"] = 24 * entries + 32 * reads + 24 * paths
            line["roofline"]["hbm_frac_of_bytes_read"] = (24 * entries + 32 * reads + 24 * paths) * its / (em_ms_mean * 1e-3) / 1e9 / peak
            line["roofline"]["us_per_iteration"] = 1e3 * em_ms_mean / its
            line["roofline"]["traffic"] = None
            line["roofline"]["traffic_source"] = None

        if not args.no_cpu_baseline and not giant:
            threads = host_threads()
            port_value, port_seconds, port_its = run_port_full(arena, threads)
            line["cpu_port"] = {"value": port_value, "unit": UNIT, "cores": threads, "kind": "port", "seconds": port_seconds,
                                "sample": "the whole arena of this rank (sparse CPU restatement, oracle/rpvg_oracle.cpp)",
                                "iterations_identical_to_gpu": bool(np.array_equal(port_its.astype(np.int64), its_all))}
            sample, _ = reference_sample(arena, its_all, threads)
            sub = arena.select(sample)
            v, s_, kind, cores, busy = run_reference_sample(sub, its_all[sample], threads)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "seconds": s_, "host_core_utilisation": busy,
                                    "sample": f"{len(sample)} of {arena.n_clusters} clusters drawn at random among those costing <= 1/(4 x threads) of the "
                                              f"budget (dense M x (N+1) x iterations)"}
            # the GPU on exactly that sample, through the host-facing call
            sp, so = pin_arena(torch, sub), pinned_outputs(torch, sub)
            gpu_ms = float(np.mean(time_e2e(torch, ctx, flush, params, sp, so, 3, 2)))
            line["same_work"] = {"clusters": len(sample), "gpu_ms": gpu_ms, "cpu_ms": 1e3 * s_, "ratio": 1e3 * s_ / gpu_ms,
                                 "cpu": f"{kind}, {cores} threads", "gpu": "rpvg_b200_batch_upload + _em + _download from pinned host buffers"}
        print(json.dumps(line))

    barrier()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
