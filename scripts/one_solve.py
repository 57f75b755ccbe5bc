"""One solve of a config (for ncu launch lists). usage: one_solve.py <config> [scale]"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from beam_slam_b200 import host
from beam_slam_b200.solver import Solver
cfg = int(sys.argv[1]); scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
g = host.HashGraph.synthetic(cfg, scale=scale); p = g.flatten(); opts = host.HashGraph.synthetic_options(cfg)
s = Solver(opts); s.set_problem(p)
for _ in range(4):
    s.set_values(p.block_values); summ, its = s.solve()
print(summ.message.decode(), summ.num_iterations, summ.device_time_in_seconds * 1e3, "ms")
