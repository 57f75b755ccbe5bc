"""GPU parity proper: the CUDA path, called through the C-ABI, against the CPU oracle on the same seeded graphs.

Tolerances (BASELINE.json north_star asks 1e-9 relative on per-iteration cost, step norm and final cost):
  * ordering and sparsity pattern: bit-exact;
  * reprojection residuals: 1e-15 (same IEEE operations in the same order); all residuals 1e-12; gradient 1e-10;
  * reduced system: 1e-9 relative to max|S|;
  * per-iteration cost, final cost, trust-region radius: 1e-9 relative at every iteration;
  * step norm: 1e-9 relative for the first step (identical linearisation point). After that the reference's
    finite-difference quaternion Jacobian (EPSILON = 1e-8, euclidean_reprojection_function.h:13-24,124-137) turns
    a 1-ulp difference in a quaternion into a ~1e-8 relative change of J_q, so two correct implementations (or the
    reference itself built with/without FMA) drift apart by ~1e-8..1e-6 in the STEP while the cost still agrees
    to 1e-9; tests/test_oracle_sensitivity.py shows the oracle doing exactly that against itself. Later steps
    are therefore held to 1e-5."""
import numpy as np
import pytest

import oracle_lib as orc
from beam_slam_b200 import cabi, host
from beam_slam_b200.builder import ProblemBuilder
from beam_slam_b200.solver import BsnlsError, Solver

pytestmark = pytest.mark.gpu

# C1, C2, C3 at the sizes BASELINE.json names, plus scaled-down ones that exercise the other code paths (C3x0.5 has a
# 600-dim reduced system: the tile-sparse Cholesky with its internal block reordering). C4 / C5 at full size:
# tests/test_gpu_parity_full.py.
CASES = [(1, 1.0), (2, 0.25), (2, 1.0), (3, 0.02), (3, 0.5), (3, 1.0)]


def _rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(scope="module", params=CASES, ids=lambda c: f"C{c[0]}x{c[1]}")
def case(request):
    cfg, scale = request.param
    g = host.HashGraph.synthetic(cfg, scale=scale)
    return cfg, g, g.flatten(), host.HashGraph.synthetic_options(cfg)


def test_ordering_and_sparsity_pattern_bit_exact(case):
    cfg, g, p, opts = case
    s = Solver(opts)
    s.set_problem(p)
    e, f = s.ordering()
    eo, fo = orc.ordering(p)
    assert np.array_equal(e, eo) and np.array_equal(f, fo)
    # Structural pattern of the reduced system in 3x3 blocks, derived independently from the flat problem the way
    # Ceres' block structure is: f-blocks co-occurring in a residual block, and f-blocks of reprojection factors that
    # share an eliminated landmark. The oracle's numeric nonzeros must be a subset of it.
    fpos = {int(b): k for k, b in enumerate(f)}
    pairs = {(k, k) for k in range(len(f))}
    for t, grp in p.groups.items():
        if t == cabi.FACTOR_REPROJECTION:
            by_lm = {}
            for q, tt, lm in grp["blocks"]:
                by_lm.setdefault(int(lm), set()).update(fpos[int(b)] for b in (q, tt) if int(b) in fpos)
            sets = by_lm.values()
        else:
            sets = [{fpos[int(b)] for b in row if int(b) in fpos} for row in grp["blocks"]]
        for st in sets:
            pairs.update((a, b) for a in st for b in st if a >= b)
    expect = np.array(sorted(pairs)).T
    r, c = s.reduced_pattern()
    assert np.array_equal(np.stack([r, c]), expect)
    S, _ = orc.reduced_system(p, 1e4, opts)
    nf = len(f)
    blk = np.abs(S.reshape(nf, 3, nf, 3)).max(axis=(1, 3)) > 0
    ro, co = np.nonzero(np.tril(blk))
    assert set(zip(ro.tolist(), co.tolist())) <= pairs


def test_evaluate_matches_oracle(case):
    cfg, g, p, opts = case
    s = Solver(opts)
    s.set_problem(p)
    cost, res, grad = s.evaluate()
    co, ro, go = orc.evaluate(p)
    assert res.shape == ro.shape and grad.shape == go.shape
    assert abs(cost - co) <= 1e-12 * co
    assert _rel(res, ro) < 1e-12
    # rows of reprojection factors: same IEEE operations in the same order on both sides
    rows = np.zeros(len(ro), bool)
    nres = np.zeros(p.num_residual_blocks, np.int64)
    for t, grp in p.groups.items():
        nres[grp["res_index"]] = cabi.FACTOR_NRES[t]
    start = np.concatenate([[0], np.cumsum(nres)])
    if cabi.FACTOR_REPROJECTION in p.groups:
        for ri in p.groups[cabi.FACTOR_REPROJECTION]["res_index"]:
            rows[start[ri]:start[ri] + 2] = True
        assert _rel(res[rows], ro[rows]) < 1e-15
    assert _rel(grad, go) < 1e-10


def test_reduced_system_matches_oracle(case):
    cfg, g, p, opts = case
    s = Solver(opts)
    s.set_problem(p)
    for radius in (1e4, 3.0):
        S, rhs = s.reduced_system(radius)
        So, rhso = orc.reduced_system(p, radius, opts)
        assert _rel(S, So) < 1e-9
        assert _rel(rhs, rhso) < 1e-9


def test_solve_matches_oracle_per_iteration(case):
    cfg, g, p, opts = case
    s = Solver(opts)
    s.set_problem(p)
    summ, its = s.solve()
    so, ito, xo = orc.solve(p, opts)
    assert summ.termination_type == so.termination_type, (summ.message, so.message)
    assert summ.message == so.message
    assert len(its) == len(ito)
    for a, b in zip(its, ito):
        assert a.iteration == b.iteration and a.step_is_successful == b.step_is_successful
        assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
        # see the module docstring (finite-difference J_q): later steps drift by ~1e-8 of the first step's length
        step_tol = 1e-9 * b.step_norm if a.iteration <= 1 else 1e-5 * b.step_norm + 1e-6 * ito[1].step_norm
        assert abs(a.step_norm - b.step_norm) <= max(step_tol, 1e-21)
        assert abs(a.trust_region_radius - b.trust_region_radius) <= 1e-9 * b.trust_region_radius
    assert abs(summ.final_cost - so.final_cost) <= 1e-9 * so.final_cost
    assert abs(summ.initial_cost - so.initial_cost) <= 1e-12 * so.initial_cost
    x = s.get_values()
    assert _rel(x, xo) < 1e-7
    assert summ.num_e_blocks == so.num_e_blocks and summ.reduced_system_dim == so.reduced_system_dim


def test_dropin_hash_graph_optimize_updates_variables_in_place():
    g = host.HashGraph.synthetic(1, scale=0.5)
    p = g.flatten()
    opts = host.HashGraph.synthetic_options(1)
    so, ito, xo = orc.solve(p, opts)
    before = g.current_values()
    summ, its = g.optimize(opts)
    after = g.current_values()
    assert summ.is_solution_usable()
    assert not np.array_equal(before, after)
    assert _rel(after, xo) < 1e-7
    assert abs(summ.final_cost - so.final_cost) <= 1e-9 * so.final_cost
    assert abs(g.evaluate() - so.final_cost) <= 1e-9 * so.final_cost


def test_relative_pose_known_answer_no_landmarks():
    """test_relative_pose_3d_stamped_constraint.cpp:124-232 on the GPU: pure pose graph (no e-blocks)."""
    b = ProblemBuilder()
    o1 = b.add_quaternion([0.952, 0.038, -0.189, 0.239]); p1 = b.add_block([1.5, -3.0, 10.0])
    o2 = b.add_quaternion([0.944, -0.128, 0.145, -0.269]); p2 = b.add_block([-1.5, 3.0, -10.0])
    I6 = np.eye(6).ravel()
    b.add_factor(cabi.FACTOR_ABSOLUTE_POSE, [p1, o1], np.concatenate([[0, 0, 0, 1, 0, 0, 0], I6]))
    b.add_factor(cabi.FACTOR_RELATIVE_POSE, [p1, o1, p2, o2], np.concatenate([[1, 0, 0, 1, 0, 0, 0], I6]))
    p = b.build()
    s = Solver()
    s.set_problem(p)
    summ, its = s.solve()
    x = s.get_values()
    assert summ.is_solution_usable()
    assert np.abs(x[4:7]).max() < 1e-5 and np.abs(x[11:14] - [1, 0, 0]).max() < 1e-5
    assert np.abs(x[0:4] - [1, 0, 0, 0]).max() < 1e-3 and np.abs(x[7:11] - [1, 0, 0, 0]).max() < 1e-3
    so, ito, xo = orc.solve(p)
    assert len(its) == len(ito)
    for a, c in zip(its, ito):
        assert abs(a.cost - c.cost) <= 1e-9 * max(abs(c.cost), 1e-30) + 1e-25


def test_imu_prior_known_answer_and_constant_blocks():
    """absolute_imu_state_3d_stamped_constraint_test.cpp:85-298 on the GPU + a held block (hash_graph.cpp:500-503)."""
    import test_oracle_golden as tg
    cov = tg._cov15()
    mean = np.array([1.0, 0, 0, 0, 1, 2, 3, 1, 2, 3, 0.1, 0.2, 0.3, 0.1, 0.2, 0.3])
    A = np.linalg.cholesky(np.linalg.inv(cov)).T
    for hold in (False, True):
        b = ProblemBuilder()
        ids = [b.add_quaternion([0.952, 0.038, -0.189, 0.239]), b.add_block([1.5, -3.0, 10.0]),
               b.add_block([1.5, -3.0, 10.0], constant=hold), b.add_block([0.15, -0.30, 1.0]), b.add_block([0.15, -0.30, 1.0])]
        b.add_factor(cabi.FACTOR_IMU_PRIOR, ids, np.concatenate([mean, A.ravel()]))
        p = b.build()
        s = Solver()
        s.set_problem(p)
        summ, its = s.solve()
        x = s.get_values()
        so, ito, xo = orc.solve(p)
        assert summ.termination_type == so.termination_type
        assert _rel(x, xo) < 1e-8
        if not hold:
            assert np.abs(x[:4] - mean[:4]).max() < 1e-3 and np.abs(x[4:] - mean[4:]).max() < 1e-5
        else:
            assert np.array_equal(x[7:10], [1.5, -3.0, 10.0])


def test_unsupported_and_invalid_inputs_fail_loudly():
    b = ProblemBuilder()
    v = b.add_block([1.0, 2.0])  # 2-vector: not on this path
    b.add_factor(cabi.FACTOR_ABSOLUTE_VEC3, [v], np.zeros(12))
    s = Solver()
    with pytest.raises(BsnlsError) as e:
        s.set_problem(b.build())
    assert e.value.code == cabi.ERR_UNSUPPORTED
    with pytest.raises(BsnlsError) as e:
        s.solve()
    assert e.value.code == cabi.ERR_STATE
    with pytest.raises(BsnlsError):
        Solver(cabi.Options.defaults(min_trust_region_radius=-1.0))


def test_failure_leaves_parameters_unchanged():
    """On FAILURE the caller's IsSolutionUsable() is false and values stay put (fixed_lag_smoother.cpp:286)."""
    b = ProblemBuilder()
    v = b.add_block([1.0, 2.0, 3.0])
    A = np.eye(3).ravel()
    b.add_factor(cabi.FACTOR_ABSOLUTE_VEC3, [v], np.concatenate([[np.nan, 0.0, 0.0], A]))
    p = b.build()
    s = Solver()
    s.set_problem(p)
    summ, its = s.solve()
    assert summ.termination_type == cabi.FAILURE and not summ.is_solution_usable()
    assert np.array_equal(s.get_values(), [1.0, 2.0, 3.0])


def test_run_to_run_determinism():
    g = host.HashGraph.synthetic(2, scale=0.25)
    p = g.flatten()
    opts = host.HashGraph.synthetic_options(2)
    out = []
    for _ in range(2):
        s = Solver(opts)
        s.set_problem(p)
        summ, its = s.solve()
        out.append((s.get_values().tobytes(), [i.cost for i in its]))
    assert out[0] == out[1]


def test_loop_closure_graph_on_the_tile_path_matches_oracle():
    """Config 4 at a tenth of its size: 200 keyframes with loop closures, reduced system 1 200. The tile-sparse Cholesky
    runs with its nested-dissection order (20 tiles, the loop-closure tiles end up in separators) and must reproduce
    the oracle's dense solve iteration by iteration."""
    g = host.HashGraph.synthetic(4, scale=0.1)
    p, opts = g.flatten(), host.HashGraph.synthetic_options(4)
    s = Solver(opts)
    s.set_problem(p)
    summ, its = s.solve()
    so, ito, xo = orc.solve(p, opts)
    assert summ.message == so.message and len(its) == len(ito)
    for a, b in zip(its, ito):
        assert a.step_is_successful == b.step_is_successful
        assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
        assert abs(a.trust_region_radius - b.trust_region_radius) <= 1e-9 * b.trust_region_radius
    assert abs(summ.final_cost - so.final_cost) <= 1e-9 * so.final_cost
    assert _rel(s.get_values(), xo) < 1e-7


@pytest.mark.parametrize("lanes", [4, 32])
def test_landmark_kernel_lane_counts(lanes, monkeypatch):
    """Four lanes per landmark on config 2 (tracks of up to 20 observations: the lanes' fourth and later observations take
    the kernel's remainder path, outside the warp-cooperative record moves) and 32 lanes: same solve as the oracle's."""
    monkeypatch.setenv("BSNLS_LM_LANES", str(lanes))
    g = host.HashGraph.synthetic(2)
    p, opts = g.flatten(), host.HashGraph.synthetic_options(2)
    s = Solver(opts)
    s.set_problem(p)
    summ, its = s.solve()
    so, ito, xo = orc.solve(p, opts)
    assert len(its) == len(ito)
    for a, b in zip(its, ito):
        assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
    assert _rel(s.get_values(), xo) < 1e-7


@pytest.mark.parametrize("order", ["graph", "band"])
def test_both_orderings_of_the_reduced_system_give_the_oracle_solve(order, monkeypatch):
    """The two candidate orders of the f-blocks (nested dissection of the block graph with tile padding; reverse
    Cuthill-McKee line + tile dissection) are only a permutation of the same factorisation: forced one at a time
    (BSNLS_ORDER) both must reproduce the oracle."""
    monkeypatch.setenv("BSNLS_ORDER", order)
    g = host.HashGraph.synthetic(4, scale=0.15)
    p, opts = g.flatten(), host.HashGraph.synthetic_options(4)
    s = Solver(opts)
    s.set_problem(p)
    summ, its = s.solve()
    so, ito, xo = orc.solve(p, opts)
    assert len(its) == len(ito)
    for a, b in zip(its, ito):
        assert abs(a.cost - b.cost) <= 1e-9 * abs(b.cost)
    assert _rel(s.get_values(), xo) < 1e-7


def test_disconnected_reduced_system_on_the_tile_path():
    """Two pose chains that share nothing (the dissection's component branch), 264 unknowns: the tile path must treat the
    components independently and agree with the oracle; one of the chains is three poses long (a region smaller than a tile)."""
    b = ProblemBuilder()
    rng = np.random.default_rng(11)
    def chain(n, origin):
        ids = []
        for k in range(n):
            q = np.array([1.0, 0.02 * k, -0.01 * k, 0.03 * k]); q /= np.linalg.norm(q)
            ids.append((b.add_block(origin + np.array([0.5 * k, 0.1 * k, 0.0]) + 0.05 * rng.normal(size=3)), b.add_quaternion(q)))
        I6 = np.eye(6).ravel()
        b.add_factor(cabi.FACTOR_ABSOLUTE_POSE, list(ids[0]), np.concatenate([origin, [1, 0, 0, 0], I6]))
        for k in range(n - 1):
            b.add_factor(cabi.FACTOR_RELATIVE_POSE, [*ids[k], *ids[k + 1]], np.concatenate([[0.5, 0.1, 0.0, 0.9995, 0.02, -0.01, 0.015], I6]),
                         loss_type=cabi.LOSS_CAUCHY, loss_param=1.0)
    chain(41, np.array([0.0, 0.0, 0.0])); chain(3, np.array([10.0, -5.0, 2.0]))
    p = b.build()
    s = Solver(); s.set_problem(p)
    summ, its = s.solve()
    so, ito, xo = orc.solve(p)
    assert summ.termination_type == so.termination_type and len(its) == len(ito)
    for a, c in zip(its, ito):
        assert abs(a.cost - c.cost) <= 1e-9 * max(abs(c.cost), 1e-30) + 1e-22
    assert _rel(s.get_values(), xo) < 1e-7


@pytest.mark.parametrize("cfg,scale", [(2, 0.25), (3, 0.02)])
def test_block_sparse_pcg_matches_dense_and_oracle(cfg, scale):
    """K7: block-sparse reduced system + block-Jacobi PCG against the dense DMMA Cholesky path and the oracle.
    The assembled system is the same arithmetic (bit-identical blocks); the PCG solve is iterative, so the
    per-iteration cost is held to 1e-8 relative (pcg_tolerance 1e-13 on |r|/|b|)."""
    g = host.HashGraph.synthetic(cfg, scale=scale)
    p = g.flatten()
    opts = host.HashGraph.synthetic_options(cfg)
    dense = Solver(opts)
    dense.set_problem(p)
    opts_s = host.HashGraph.synthetic_options(cfg)
    opts_s.linear_solver = cabi.LINEAR_ITERATIVE_SCHUR
    opts_s.pcg_max_iterations = 5000
    sp = Solver(opts_s)
    sp.set_problem(p)
    Sd, rd = dense.reduced_system(1e4)
    Ss, rs = sp.reduced_system(1e4)
    assert np.array_equal(np.tril(Sd), np.tril(Ss)) and np.array_equal(rd, rs)
    summ, its = sp.solve()
    so, ito, xo = orc.solve(p, opts)
    assert summ.linear_solver_used == cabi.LINEAR_ITERATIVE_SCHUR
    assert summ.termination_type == so.termination_type and len(its) == len(ito)
    assert all(i.linear_solver_iterations > 0 for i in its[1:])
    for a, b in zip(its, ito):
        assert abs(a.cost - b.cost) <= 1e-8 * abs(b.cost)
    assert abs(summ.final_cost - so.final_cost) <= 1e-9 * so.final_cost
    r, c = sp.reduced_pattern()
    assert summ.reduced_system_nnz_blocks == len(r)
