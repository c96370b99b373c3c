"""bench.py --config c1|c3|c4|c5 (part of bench.py, kept in a file of its own): the model classes of the other BASELINE.json configurations through the batched driver.

A step = one ``estimateBatch`` call of the drop-in class over the whole (per-rank share of the) configuration: flatten ->
H2D -> kernels -> D2H -> fill of every PathClusterEstimates, i.e. the inference stage the reference times at
rpvg:src/main.cpp:1090-1091 from the point the boundary inputs exist. The call takes and returns host
objects, so the number is end to end by construction. ``value`` = read pairs (sum of the read-class counts) per second of
that call, max over ranks; ``inference_stage_wall_ms`` is the wall time itself. N > 1: the clusters of the one job are dealt
to the ranks by LPT (fixed total work: "scaling": "strong"), no collective on the data path.

--impl reference: the reference's own classes (oracle/_ref) with all host threads, rpvg's largest-first dynamic loop, on a
bounded random sample of the clusters (those whose dense cost estimate stays below a cap), chunk after chunk until about
15 s have passed; the b200 arm runs the identical sample as ``same_work``.

c4 additionally times the pair log-likelihood kernel alone on resident matrices (rpvg_b200_grouped_raw +
rpvg_b200_grouped_bounded_pairs) against the fp64 log throughput this GPU sustains (rpvg_b200_log_peak): the roofline of
the haplotype models is log-bound, not HBM-bound.
"""

from __future__ import annotations

import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def build_config(name, scale):
    from rpvg_b200 import synth
    from tests.harness import binding as loader
    if name == "c1":
        arena = synth.example_arena(n_reads=int(100_000 * scale))
        cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
        what = "config1: example-structured haplotype-transcripts (36,120 paths, 2,177 transcripts, 1,729 haplotype ids, 100k read pairs; synthetic reads, seed 20240602)"
    elif name == "c3":
        n_transcripts = max(50, int(2000 * scale))
        arena = synth.thousand_genomes_arena(n_transcripts=n_transcripts, reads_per_transcript=300)
        cfg = loader.make_config(loader.MODEL_HAPLOTYPE_TRANSCRIPTS, group_size=2)
        what = (f"config3: haplotype-transcripts at 1000GP scale, 5,008 haplotype ids, {n_transcripts} of the 100,000 transcripts "
                f"(knob; 300 read pairs per transcript; seed 20240604)")
    elif name == "c4":
        arena = synth.haplotypes_arena(n_clusters=max(10, int(200 * scale)), n_max=2000)
        cfg = loader.make_config(loader.MODEL_HAPLOTYPES, group_size=2)
        what = "config4: haplotypes model, diploid bounded posteriors, 200 clusters of 2-2,000 haplotype-specific paths (seed 20240605)"
    else:
        arena = synth.strains_arena(n_paths=int(200_000 * scale), n_reads=int(1_000_000 * scale))
        cfg = loader.make_config(loader.MODEL_STRAINS)
        what = "config5: strains model (minimum path cover + EM), config 2's matrices with 30% of the paths made redundant (seed 20240606)"
    return arena, cfg, what


def dense_cost(arena, model_is_nested):
    """What the reference pays for a cluster, roughly: pairs x read classes for the haplotype models (its columns are the
    haplotype groups: at most paths x transcripts-in-cluster combinations), M x N x ~1000 iterations for the strains EM."""
    reads = np.diff(arena.read_offset.astype(np.int64)).astype(np.float64)
    paths = np.diff(arena.path_offset.astype(np.int64)).astype(np.float64)
    if arena.group_id is None:
        return reads * (paths + 1) * (paths + 200.0)
    groups = paths.copy()
    if model_is_nested:
        for c in range(arena.n_clusters):
            p0, p1 = int(arena.path_offset[c]), int(arena.path_offset[c + 1])
            transcripts = len(np.unique(arena.group_id[p0:p1]))
            if transcripts > 1:
                groups[c] = min(5008.0, paths[c] * 8.0 ** (transcripts - 1))
    return groups * (groups + 1) / 2 * np.maximum(reads, 1)


def run(args, rank, local_rank, world, ClockSampler, host_threads):
    sys.path.insert(0, ROOT)
    from tests.harness import binding as loader   # marshalling for the estimator harness (this repository's classes)
    from rpvg_b200 import multi_gpu

    arena, cfg, what = build_config(args.config, args.scale)
    nested = cfg.model == loader.MODEL_HAPLOTYPE_TRANSCRIPTS
    config = {"workload": what if args.scale == 1.0 else f"DEBUG scale={args.scale}: {what}", "clusters": arena.n_clusters, "paths": arena.n_paths,
              "read_classes": arena.n_reads, "sharding": f"clusters dealt to {world} rank(s) by LPT on the model's cost (pairs x read classes, or path non-zeros), no collective",
              "l2": "inputs come from host memory every step (nothing stays resident between steps)"}
    metric, unit = "inference_stage_reads_per_second", "reads/s"
    cost = dense_cost(arena, nested)

    def sample_clusters(threads, seed=3, cap=4e9):
        eligible = np.nonzero((cost <= cap) & (np.diff(arena.read_offset.astype(np.int64)) > 0))[0]
        return np.random.default_rng(seed).permutation(eligible)

    def reads_of(ids):
        return float(sum(int(arena.count[int(arena.read_offset[c]):int(arena.read_offset[c + 1])].sum()) for c in ids))

    # ------------------------------------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        from oracle import loader as oracle_loader   # the reference arm: rpvg's own classes
        threads = host_threads()
        ref = oracle_loader.reference()
        order = sample_clusters(threads)
        chunk = 4 * threads
        sample, seconds = [], 0.0
        for k in range(0, len(order), chunk):
            ids = [int(c) for c in order[k:k + chunk]]
            s, res = loader.harness_estimate_batch(ref, cfg, arena, ids, rng_seed=1, num_threads=threads)
            sample.extend(ids)
            seconds += s
            if seconds > 15.0:
                break
        value = reads_of(sample) / seconds
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": 1e3 * seconds,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config, "impl": "reference",
                "cpu_baseline": {"value": value, "unit": unit, "cores": threads, "kind": "reference",
                                 "sample": f"{len(sample)} of {arena.n_clusters} clusters (random among those below the dense-cost cap), chunks of {chunk}"},
                "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------------------------------------------ B200 arm
    import torch
    import torch.distributed as dist
    from rpvg_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")

    os.environ["RPVG_B200_DEVICE"] = str(local_rank)
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # torchrun sets OMP_NUM_THREADS=1 for its workers: the host stages of estimateBatch (flatten, haplotype groups, plans)
    # are OpenMP loops, every rank gets its share of the cores (read when the harness library brings in the OpenMP runtime)
    if world > 1:
        os.environ["OMP_NUM_THREADS"] = str(max(1, host_threads() // world))

    harness = loader.load_harness(os.path.join(ROOT, "tests", "_build", "libharness_b200.so"))
    harness.harness_set_threads(max(1, host_threads() // world))
    # clusters dealt by the cost of the model: pairs x read classes for the haplotype models, path non-zeros otherwise
    nnz = np.diff(arena.row_ptr[arena.read_offset.astype(np.int64)].astype(np.int64)).astype(np.float64) + 1
    weight = cost + nnz if (nested or cfg.model == loader.MODEL_HAPLOTYPES) else nnz
    mine = [int(c) for c in multi_gpu.lpt_partition(weight, world)[rank]]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = capi.Context(local_rank)
    sampler = ClockSampler(local_rank)

    for _ in range(args.warmup):
        loader.harness_estimate_batched(harness, cfg, arena, mine, rng_seed=1)

    barrier()
    sampler.start()
    launches_before = ctx.launches
    seconds = []
    for _ in range(args.steps):
        s, _ = loader.harness_estimate_batched(harness, cfg, arena, mine, rng_seed=1)
        seconds.append(s)
    barrier()
    clocks = sampler.stop()

    total = torch.tensor([float(np.sum(seconds))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.MAX)
    max_seconds = float(total.cpu()[0])

    if rank == 0:
        reads = reads_of(range(arena.n_clusters))
        value = reads * args.steps / max_seconds
        h2d = arena.input_bytes()
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * max_seconds / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config, "clocks": clocks, "inference_stage_wall_ms": 1e3 * max_seconds / args.steps,
                "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8 * arena.n_paths,
                        "ms_per_step": 1e3 * max_seconds / args.steps,
                        "note": "estimateBatch takes and fills host objects: the value is end to end by construction; h2d bytes = the expanded arena (the "
                                "haplotypes model sends its read records unexpanded, about a third of that)"},
                "gpu_launches": None}

        if args.config == "c4":
            line["roofline"] = pair_kernel_roofline(ctx, arena)

        from oracle import loader as oracle_loader   # cpu_baseline leg only
        if not args.no_cpu_baseline and oracle_loader.reference_available():
            threads = host_threads()
            ref = oracle_loader.reference()
            order = sample_clusters(threads)
            chunk = 4 * threads
            sample, cpu_s = [], 0.0
            for k in range(0, len(order), chunk):
                ids = [int(c) for c in order[k:k + chunk]]
                s, _ = loader.harness_estimate_batch(ref, cfg, arena, ids, rng_seed=1, num_threads=threads)
                sample.extend(ids)
                cpu_s += s
                if cpu_s > 15.0:
                    break
            loader.harness_estimate_batched(harness, cfg, arena, sample, rng_seed=1)
            gpu_s = min(loader.harness_estimate_batched(harness, cfg, arena, sample, rng_seed=1)[0] for _ in range(3))
            line["cpu_baseline"] = {"value": reads_of(sample) / cpu_s, "unit": unit, "cores": threads, "kind": "reference", "seconds": cpu_s,
                                    "sample": f"{len(sample)} of {arena.n_clusters} clusters (random among those below the dense-cost cap)"}
            line["same_work"] = {"clusters": len(sample), "gpu_ms": 1e3 * gpu_s, "cpu_ms": 1e3 * cpu_s, "ratio": cpu_s / gpu_s,
                                 "cpu": f"reference classes, {threads} threads", "gpu": "estimateBatch of the drop-in class"}
        print(json.dumps(line))

    barrier()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def pair_kernel_roofline(ctx, arena):
    """The pair log-likelihood kernel alone, on the raw matrices of every cluster resident in HBM: terms (pair x read
    class) per second. A term costs the kernel two fp64 instructions (the sum and the running product it goes into), so
    half of the measured fp64 rate is its roofline; the rate a log per term would allow (rpvg_b200_log_peak), which bounds
    the reference's formulation, is reported beside it."""
    from rpvg_b200 import capi
    lib = capi.lib()
    lib.rpvg_b200_log_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    lib.rpvg_b200_fp64_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    lib.rpvg_b200_grouped_raw.argtypes = [C.c_void_p, C.POINTER(capi.Clusters), C.POINTER(C.c_void_p)]
    lib.rpvg_b200_grouped_bounded_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double] + [C.POINTER(C.c_void_p)] * 3
    lib.rpvg_b200_grouped_pair_stats.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.rpvg_b200_grouped_free.argtypes = [C.c_void_p, C.c_void_p]

    peak = C.c_double()
    ctx._check(lib.rpvg_b200_log_peak(ctx._h, C.byref(peak)))
    fma_peak = C.c_double()
    ctx._check(lib.rpvg_b200_fp64_peak(ctx._h, C.byref(fma_peak)))

    keep = [c for c in range(arena.n_clusters) if arena.read_offset[c + 1] > arena.read_offset[c]]
    sub = arena.select(keep)
    cl = capi.clusters_struct(sub)
    handle = C.c_void_p()
    ctx._check(lib.rpvg_b200_grouped_raw(ctx._h, C.byref(cl), C.byref(handle)))

    group_offset = np.ascontiguousarray(sub.path_offset, dtype=np.uint64)
    order = np.concatenate([np.arange(int(sub.path_offset[c + 1] - sub.path_offset[c]), dtype=np.uint32) for c in range(sub.n_clusters)])
    log_freq = np.zeros(sub.n_paths, dtype=np.float64)
    outs = [C.c_void_p(), C.c_void_p(), C.c_void_p()]
    for _ in range(3):   # first call = warm-up; the statistics add up over the calls
        ctx._check(lib.rpvg_b200_grouped_bounded_pairs(ctx._h, handle, group_offset.ctypes.data, order.ctypes.data, log_freq.ctypes.data, float(np.log(1e-8)),
                                                       C.byref(outs[0]), C.byref(outs[1]), C.byref(outs[2])))
        if _ == 0:
            ms0, logs0 = C.c_double(), C.c_double()
            lib.rpvg_b200_grouped_pair_stats(handle, C.byref(ms0), C.byref(logs0))
    ms, logs = C.c_double(), C.c_double()
    lib.rpvg_b200_grouped_pair_stats(handle, C.byref(ms), C.byref(logs))
    lib.rpvg_b200_grouped_free(ctx._h, handle)

    kernel_ms = (ms.value - ms0.value) / 2
    evaluations = (logs.value - logs0.value) / 2
    achieved = evaluations / (kernel_ms * 1e-3)
    term_peak = fma_peak.value / 2
    return {"bound": "fp64 pipe", "kernel": "pairTileKernel (+ pairCombineKernel)", "achieved": achieved / 1e9, "peak": term_peak / 1e9, "unit": "Gterm/s",
            "frac": achieved / term_peak, "traffic": None,
            "peak_source": "measured in this run: rpvg_b200_fp64_peak (independent fp64 fma chains, no memory traffic) / 2 fp64 instructions per term",
            "terms_per_launch": evaluations, "kernel_ms": kernel_ms,
            "log_bound": {"peak": peak.value / 1e9, "unit": "Glog/s", "times": achieved / peak.value,
                          "meaning": "rate a log per term would allow (rpvg_b200_log_peak): the bound of the reference's formulation, which this kernel replaces by "
                                     "one log per pair and row range"},
            "algorithmic": "pairs x read classes: sum_c G_c (G_c + 1) / 2 x M_c terms (SURVEY.md §8d)"}
