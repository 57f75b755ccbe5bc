"""Landmark-sharded solve across GPUs (NCCL all-reduce of the reduced camera system) against the single-GPU solve.
Needs >= 2 GPUs; skipped on a single-GPU box (the host-side partition logic is covered on CPU by
tests/test_sharding_cpu.py)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, cfg, scale, linear_solver, out):
    import torch.distributed as dist
    from beam_slam_b200 import cabi, host, sharding
    from beam_slam_b200.solver import Solver
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    p = host.HashGraph.synthetic(cfg, scale=scale).flatten()
    opts = host.HashGraph.synthetic_options(cfg)
    opts.linear_solver = linear_solver
    opts.device = rank
    s = sharding.create_sharded_solver(opts)
    s.set_problem(p)                      # every rank passes the full problem; the library keeps its shard
    S, rhs = s.reduced_system(1e4)
    summ, its = s.solve()
    x = s.get_values()
    shards = [None] * world
    dist.all_gather_object(shards, s.shard().tolist())
    if rank == 0:
        e_blocks, _ = s.ordering()
        out["partition"] = (sorted(sum(shards, [])) == list(range(len(e_blocks))), [len(v) for v in shards])
        ref = Solver(opts)
        ref.set_problem(p)
        S1, rhs1 = ref.reduced_system(1e4)
        s1, it1 = ref.solve()
        x1 = ref.get_values()
        out["S"] = float(np.abs(S - S1).max() / np.abs(S1).max())
        out["rhs"] = float(np.abs(rhs - rhs1).max() / np.abs(rhs1).max())
        out["n"] = (len(its), len(it1))
        out["cost"] = max(abs(a.cost - b.cost) / b.cost for a, b in zip(its, it1))
        out["final"] = abs(summ.final_cost - s1.final_cost) / s1.final_cost
        out["x"] = float(np.abs(x - x1).max() / np.abs(x1).max())
        out["term"] = (summ.termination_type, s1.termination_type)
    dist.barrier()
    s.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("cfg,scale,ls", [(3, 0.05, 0), (3, 0.05, 1), (4, 0.12, 0)])   # the last one is large enough for the on-chip chunk path
def test_sharded_solve_matches_single_gpu(cfg, scale, ls):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), cfg, scale, ls, out), nprocs=world, join=True)
    assert out["partition"][0] and min(out["partition"][1]) > 0   # the shards partition the landmarks (locality order, balanced by observations)
    assert out["S"] < 1e-12 and out["rhs"] < 1e-12           # same blocks, different summation split
    assert out["n"][0] == out["n"][1] and out["term"][0] == out["term"][1]
    assert out["cost"] < 1e-9 and out["final"] < 1e-9
    assert out["x"] < 1e-6
