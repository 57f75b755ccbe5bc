"""Per-kernel split of one linearisation pass of a config (direct launches, CUDA events): python scripts/pass_breakdown.py <config> [scale]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from beam_slam_b200 import host
from beam_slam_b200.solver import Solver
cfg = int(sys.argv[1]); scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
g = host.HashGraph.synthetic(cfg, scale=scale); p = g.flatten(); o = host.HashGraph.synthetic_options(cfg)
s = Solver(o); s.set_problem(p)
for _ in range(3): s.run_pass(1e4)
tot = []
for _ in range(10): s.run_pass(1e4); tot.append(s.last_kernel_ms()[0])
s.enable_breakdown(True)
parts = []
for _ in range(5): s.run_pass(1e4); parts.append(s.last_breakdown_ms())
print("pass (graph) ms: median %.4f min %.4f" % (np.median(tot), min(tot)))
print({k: round(float(np.mean([q[k] for q in parts])), 4) for k in parts[0]})
