"""Host logic of the multi-GPU path on the CPU: LPT partition and a world_size-2 gloo run whose per-rank compute is the
CPU oracle (this is a test of sharding / gathering / reassembly, not of the kernels)."""

import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_lpt_partition_balances_and_covers():
    from rpvg_b200 import multi_gpu, synth
    arena = synth.transcripts_arena(seed=3, n_paths=20000, n_reads=100000, n_max=2000)
    costs = multi_gpu.cluster_costs(arena)
    for parts_n in (1, 2, 4, 8):
        parts = multi_gpu.lpt_partition(costs, parts_n)
        flat = np.sort(np.concatenate(parts))
        assert np.array_equal(flat, np.arange(arena.n_clusters)), "every cluster exactly once"
        loads = np.array([costs[p].sum() for p in parts])
        # LPT bound: max load <= mean + largest single cost
        assert loads.max() <= loads.mean() + costs.max() + 1e-9
        if parts_n > 1 and costs.max() < loads.mean():
            assert loads.max() / loads.mean() < 1.05
    assert [list(p) for p in multi_gpu.lpt_partition(costs, 4)] == [list(p) for p in multi_gpu.lpt_partition(costs, 4)]


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import loader
    from rpvg_b200 import multi_gpu, synth

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    arena = synth.transcripts_arena(seed=21, n_paths=1500, n_reads=8000, n_max=200)

    def compute(sub):
        ab, noise, total, its, _ = loader.port_em_abundances(sub, num_threads=1)
        return ab, noise, total, its

    def gather(obj):
        out = [None] * world if rank == 0 else None
        dist.gather_object(obj, out, dst=0)
        return out

    res = multi_gpu.run_sharded(arena, compute, rank, world, gather)
    if rank == 0:
        np.savez(out_path, abundances=res[0], noise=res[1], total=res[2], its=res[3])
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_run_equals_single_process(tmp_path):
    from oracle import loader
    from rpvg_b200 import synth

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]

    out_path = str(tmp_path / "sharded.npz")
    mp.spawn(_worker, args=(2, port, out_path), nprocs=2, join=True)

    arena = synth.transcripts_arena(seed=21, n_paths=1500, n_reads=8000, n_max=200)
    ab, noise, total, its, _ = loader.port_em_abundances(arena)
    z = np.load(out_path)
    assert np.array_equal(z["abundances"], ab) and np.array_equal(z["noise"], noise)
    assert np.array_equal(z["total"], total) and np.array_equal(z["its"], its)
