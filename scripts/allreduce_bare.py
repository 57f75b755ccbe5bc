"""Bare NCCL all-reduce of a buffer the size of the sharded solve's exchange (config 4: 6.7 MB of float64), for comparison with
the `exchange` entry of bench.py's kernel split (pack + all-reduce + unpack). torchrun --nproc-per-node N scripts/allreduce_bare.py [MB]"""
import os, sys
import torch, torch.distributed as dist
mb = float(sys.argv[1]) if len(sys.argv) > 1 else 6.7
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
x = torch.ones(int(mb * 1e6 / 8), dtype=torch.float64, device="cuda")
for n in (1, 8, len(x)):
    y = x[:n]
    for _ in range(10): dist.all_reduce(y)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20): dist.all_reduce(y)
    g.replay(); torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    if dist.get_rank() == 0: print(f"all_reduce of {n * 8 / 1e6:.3f} MB float64 over {dist.get_world_size()} GPUs: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us (graph replay of 20)", flush=True)
dist.destroy_process_group()
