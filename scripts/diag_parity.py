"""Per-iteration relative differences GPU vs oracle (diagnostic; run on the GPU box)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as orc
from beam_slam_b200 import host, cabi
from beam_slam_b200.solver import Solver

for cfg, scale in ((1, 1.0), (2, 0.25), (3, 0.02)):
    g = host.HashGraph.synthetic(cfg, scale=scale); p = g.flatten(); opts = host.HashGraph.synthetic_options(cfg)
    s = Solver(opts); s.set_problem(p)
    S, rhs = s.reduced_system(1e4); So, rhso = orc.reduced_system(p, 1e4, opts)
    d = np.abs(S - So); print(f"C{cfg}: S max abs diff {d.max():.3e} (max|S| {np.abs(So).max():.3e}); elementwise rel (|So|>1e-6*max): "
          f"{(d / np.maximum(np.abs(So), 1e-300))[np.abs(So) > 1e-6 * np.abs(So).max()].max():.3e}; rhs rel {np.abs(rhs - rhso).max() / np.abs(rhso).max():.3e}")
    cost, res, grad = s.evaluate(); co, ro, go = orc.evaluate(p)
    nrep = 2 * p.count(cabi.FACTOR_REPROJECTION)
    print(f"   cost rel {abs(cost-co)/co:.3e}  res maxabs diff {np.abs(res-ro).max():.3e}  grad rel {np.abs(grad-go).max()/np.abs(go).max():.3e}")
    s.set_values(p.block_values)
    summ, its = s.solve(); so, ito, xo = orc.solve(p, opts)
    print(f"   gpu: {summ.message.decode()} {summ.num_iterations} its dev {summ.device_time_in_seconds*1e3:.3f} ms | oracle: {so.message.decode()} {so.num_iterations} its {so.total_time_in_seconds*1e3:.1f} ms")
    for a, b in zip(its, ito):
        rel = lambda u, v: abs(u - v) / max(abs(v), 1e-300)
        print(f"   it {a.iteration}: cost {a.cost:.12e} rel {rel(a.cost,b.cost):.2e} | step {rel(a.step_norm,b.step_norm):.2e} | gmax {rel(a.gradient_max_norm,b.gradient_max_norm):.2e} | rho {rel(a.relative_decrease,b.relative_decrease):.2e} | radius {rel(a.trust_region_radius,b.trust_region_radius):.2e} ok {a.step_is_successful}/{b.step_is_successful}")
    x = s.get_values(); print(f"   x rel {np.abs(x-xo).max()/np.abs(xo).max():.3e}")
