"""Why later step norms are only comparable to ~1e-6: the reference's reprojection cost function differentiates
the quaternion by a forward difference with EPSILON = 1e-8 (euclidean_reprojection_function.h:13-24,124-137),
which divides rounding noise by 1e-8. Perturbing ONE quaternion component of the start point by one ulp makes the
oracle disagree with itself by far more than 1e-9 in the step norm of later iterations, while costs still agree
to 1e-9. CPU only."""
import numpy as np

import oracle_lib as orc
from beam_slam_b200 import cabi, host


def test_one_ulp_changes_later_step_norms_but_not_costs():
    g = host.HashGraph.synthetic(1, scale=0.5)
    p = g.flatten()
    opts = host.HashGraph.synthetic_options(1)
    s0, it0, x0 = orc.solve(p, opts)
    v = p.block_values.copy()
    qb = int(np.flatnonzero((p.block_manifold == cabi.MANIFOLD_QUATERNION) & (p.block_constant == 0))[3])
    k = p.offsets[qb] + 1
    v[k] = np.nextafter(v[k], np.inf)
    s1, it1, x1 = orc.solve(p.with_values(v), opts)
    assert len(it0) == len(it1)
    worst_step = max(abs(a.step_norm - b.step_norm) / b.step_norm for a, b in zip(it0[2:], it1[2:]))
    worst_cost = max(abs(a.cost - b.cost) / b.cost for a, b in zip(it0, it1))
    assert worst_cost < 1e-9
    assert worst_step > 1e-9          # the 1e-9 step-norm criterion cannot hold beyond the first step ...
    assert worst_step < 1e-4          # ... but the drift stays small (1e-5 .. 1e-6 of the step, depending on summation order)
