"""Whole-solve latency of the sliding-window configs (C1, C2) on the GPU: median device / wall time of repeated solves."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from beam_slam_b200 import host
from beam_slam_b200.solver import Solver
for cfg in (1, 2, 3):
    g = host.HashGraph.synthetic(cfg); p = g.flatten(); opts = host.HashGraph.synthetic_options(cfg)
    s = Solver(opts); s.set_problem(p)
    dev, wall = [], []
    for i in range(12):
        s.set_values(p.block_values)
        t = time.perf_counter(); summ, its = s.solve(); wall.append(time.perf_counter() - t); dev.append(summ.device_time_in_seconds)
    print(f"C{cfg}: {g.info['num_observations']} obs, {summ.num_iterations} records, {summ.message.decode()} | device ms median {np.median(dev[2:])*1e3:.3f} min {min(dev[2:])*1e3:.3f} | "
          f"wall ms median {np.median(wall[2:])*1e3:.3f} | first call wall {wall[0]*1e3:.1f} ms | launches/solve {s.launch_count()/12:.0f} | final cost {summ.final_cost:.9e}")
