import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from beam_slam_b200 import host
from beam_slam_b200.solver import Solver
g = host.HashGraph.synthetic(1); p = g.flatten(); opts = host.HashGraph.synthetic_options(1)
s = Solver(opts); s.set_problem(p)
dev = []
for i in range(12):
    s.set_values(p.block_values); summ, its = s.solve(); dev.append(summ.device_time_in_seconds)
print(os.environ.get("BSNLS_TINY_THREADS"), "C1 device ms median %.3f" % (np.median(dev[2:]) * 1e3), summ.num_iterations, "%.9e" % summ.final_cost)
