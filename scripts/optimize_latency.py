"""Host wall-clock split of GpuHashGraph::optimize for a config: first call, repeated calls on the unchanged graph, and (config 1)
the smoother cycle. BSNLS_DEBUG=1 adds the phases of bsnls_set_problem.  python scripts/optimize_latency.py <config> [scale]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from beam_slam_b200 import host
cfg = int(sys.argv[1]); scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
g = host.HashGraph.synthetic(cfg, scale=scale); o = host.HashGraph.synthetic_options(cfg)
x0 = g.flatten().block_values.copy()
for i in range(4):
    g.set_values(x0)
    t = time.perf_counter(); s, its = g.optimize(o); w = time.perf_counter() - t
    print(f"config {cfg} call {i}: wall {1e3*w:.3f} ms, device {1e3*s.device_time_in_seconds:.3f} ms, {len(its)} iterations,", {k: round(1e3*v, 3) if k.endswith('_s') else v for k, v in g.last_timing().items()})
if cfg == 1:
    r = host.SmootherReplay(20, 10)
    for i in range(8):
        u = r.advance(); t = time.perf_counter(); s, its, tm = r.optimize(o); w = time.perf_counter() - t
        print(f"cycle {i}: update {1e3*u:.3f} ms, optimize wall {1e3*w:.3f} ms, device {1e3*s.device_time_in_seconds:.3f} ms, {len(its)} its,", {k: round(1e3*v, 3) if k.endswith('_s') else v for k, v in tm.items()})
