"""The drop-in call path: GpuHashGraph::optimize / optimizeFor (host/gpu_hash_graph.cpp) as the fixed-lag smoother and
the refinement code call it (bs_optimizers/src/fixed_lag_smoother.cpp:281, bs_models/src/visual_odometry.cpp:334)."""
import numpy as np
import pytest

import oracle_lib as orc
from beam_slam_b200 import cabi, host

pytestmark = pytest.mark.gpu


def test_smoother_cycles_match_oracle_and_rebuild_only_on_structure_change():
    """Fixed-lag smoother pattern (fixed_lag_smoother.cpp:220-281): every cycle one keyframe enters, the oldest is
    pseudo-marginalised, the graph is optimised. The structure changes every cycle, so the device problem is rebuilt
    (through the cached flattener and the arena); a second optimize() of the unchanged graph only moves values."""
    r = host.SmootherReplay(16, 10)
    opts = host.HashGraph.synthetic_options(1)
    for cycle in range(5):
        r.advance()
        p = r.flatten()
        so, ito, xo = orc.solve(p, opts)
        summ, its, timing = r.optimize(opts)
        assert timing["structure_reused"] == 0.0
        assert summ.is_solution_usable() and summ.message == so.message, (cycle, summ.message, so.message)
        assert len(its) == len(ito)
        for a, b in zip(its, ito):
            assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost), (cycle, a.iteration)
        x = r.current_values(p.num_values)          # written back in place, createProblem order
        assert np.abs(x - xo).max() <= 1e-7 * np.abs(xo).max()
        summ2, its2, timing2 = r.optimize(opts)      # nothing changed but the values
        assert timing2["structure_reused"] == 1.0
        assert abs(its2[0].cost - summ.final_cost) <= 1e-12 * summ.final_cost


def test_repeated_optimize_of_an_unchanged_graph_reuses_the_device_problem():
    g = host.HashGraph.synthetic(3, scale=0.1)
    opts = host.HashGraph.synthetic_options(3)
    x0 = g.flatten().block_values.copy()
    s1, it1 = g.optimize(opts)
    assert g.last_timing()["structure_reused"] == 0.0
    g.set_values(x0)
    s2, it2 = g.optimize(opts)
    assert g.last_timing()["structure_reused"] == 1.0
    assert [i.cost for i in it1] == [i.cost for i in it2]          # same arithmetic, bit for bit
    assert s1.final_cost == s2.final_cost


def test_optimize_for_keeps_the_time_budget():
    """HashGraph::optimizeFor (hash_graph.cpp:428-444): createProblem time is charged against the budget. A budget
    that is already spent stops after iteration 0 (Ceres checks the clock when it finalises an iteration): usable,
    NO_CONVERGENCE, values untouched. A generous budget changes nothing."""
    g = host.HashGraph.synthetic(2, scale=0.5)
    opts = host.HashGraph.synthetic_options(2)
    before = g.flatten().block_values.copy()
    s0, it0 = g.optimize(opts, max_time_s=0.0)
    assert s0.termination_type == cabi.NO_CONVERGENCE and s0.is_solution_usable()
    assert s0.message == b"Maximum solver time reached." and len(it0) == 1
    assert np.array_equal(g.current_values(), before)
    sa, ita = g.optimize(opts, max_time_s=60.0)
    g.set_values(before)
    sb, itb = g.optimize(opts)
    assert [i.cost for i in ita] == [i.cost for i in itb] and sa.message == sb.message
