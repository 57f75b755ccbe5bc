"""PCG vs dense on the GPU for the refinement configs (diagnostic; run on the GPU box)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from beam_slam_b200 import host, cabi
from beam_slam_b200.solver import Solver

for cfg, scale in ((3, 1.0), (4, float(sys.argv[1]) if len(sys.argv) > 1 else 1.0)):
    g = host.HashGraph.synthetic(cfg, scale=scale); p = g.flatten()
    print(f"C{cfg} x{scale}: {g.info}")
    res = {}
    for name, ls, tol in (("dense", cabi.LINEAR_DENSE_SCHUR, 0), ("pcg1e-13", cabi.LINEAR_ITERATIVE_SCHUR, 1e-13), ("pcg1e-10", cabi.LINEAR_ITERATIVE_SCHUR, 1e-10), ("pcg1e-6", cabi.LINEAR_ITERATIVE_SCHUR, 1e-6)):
        o = host.HashGraph.synthetic_options(cfg); o.linear_solver = ls; o.pcg_max_iterations = 20000
        if tol: o.pcg_tolerance = tol
        s = Solver(o); t = time.time(); s.set_problem(p); ts = time.time() - t
        summ, its = s.solve()
        res[name] = its
        print(f"  {name:9s} setup {ts:.2f}s: {summ.message.decode()} its {summ.num_iterations} dev {summ.device_time_in_seconds*1e3:.1f} ms final {summ.final_cost:.12e} "
              f"cg iters {[i.linear_solver_iterations for i in its]} nnzb {summ.reduced_system_nnz_blocks}")
        s.close()
    for name in ("pcg1e-13", "pcg1e-10", "pcg1e-6"):
        d = [abs(a.cost - b.cost) / b.cost for a, b in zip(res[name], res["dense"])]
        print(f"  {name} vs dense: max rel cost diff {max(d):.2e}; step {max(abs(a.step_norm-b.step_norm)/max(b.step_norm,1e-300) for a,b in zip(res[name][1:], res['dense'][1:])):.2e}")
