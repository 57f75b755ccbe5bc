"""Sharded solve with the iterative linear solver: block rows of S partitioned over the ranks (default) against the same
conjugate gradients replicated on every rank (BSNLS_PCG_REPLICATED=1) and against the replicated tile Cholesky.
torchrun --nproc-per-node N scripts/sharded_pcg.py <config> [scale]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from beam_slam_b200 import cabi, host, sharding
cfg = int(sys.argv[1]); scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
p = host.HashGraph.synthetic(cfg, scale=scale).flatten()
for name, ls, env in (("tile Cholesky, replicated", cabi.LINEAR_DENSE_SCHUR, None), ("PCG, rows partitioned", cabi.LINEAR_ITERATIVE_SCHUR, None),
                      ("PCG, replicated", cabi.LINEAR_ITERATIVE_SCHUR, "1")):
    if env: os.environ["BSNLS_PCG_REPLICATED"] = env
    else: os.environ.pop("BSNLS_PCG_REPLICATED", None)
    o = host.HashGraph.synthetic_options(cfg); o.linear_solver = ls; o.device = int(os.environ["LOCAL_RANK"])
    s = sharding.create_sharded_solver(o); s.set_problem(p)
    for _ in range(2):
        s.set_values(p.block_values); summ, its = s.solve()
    if rank == 0:
        print(f"{name:28s}: {summ.num_iterations} LM iterations, final cost {summ.final_cost:.12e}, device {1e3 * summ.device_time_in_seconds:.2f} ms, "
              f"linear iterations per LM step {[it.linear_solver_iterations for it in its][1:]}", flush=True)
    dist.barrier(); s.close()
dist.destroy_process_group()
