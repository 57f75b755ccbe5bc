"""fuse_constraints::MarginalConstraint as a factor of the solve: the linear prior that
marginalizeVariables leaves behind (dependencies/fuse/fuse_constraints/src/marginalize_variables.cpp:166-258), evaluated as
MarginalCostFunction does (src/marginal_cost_function.cpp:58-113)."""
import numpy as np
import pytest

import oracle_lib as orc
from beam_slam_b200 import cabi
from beam_slam_b200.builder import ProblemBuilder


def _quat(v):
    v = np.asarray(v, float)
    return v / np.linalg.norm(v)


def _minus(xbar, x):
    """Orientation3DLocalParameterization::Minus (fuse_variables/orientation_3d_stamped.h:117-128)."""
    w1, x1, y1, z1 = xbar[0], -xbar[1], -xbar[2], -xbar[3]
    w2, x2, y2, z2 = x
    q = np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                  w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])
    s = np.linalg.norm(q[1:])
    two_theta = 2.0 * (np.arctan2(-s, -q[0]) if q[0] < 0 else np.arctan2(s, q[0]))
    return q[1:] * (two_theta / s)


def _problem(seed=3, rows=9, with_chain=True):
    rng = np.random.default_rng(seed)
    b = ProblemBuilder()
    q1, q1bar = _quat([0.93, 0.1, 0.2, 0.3]), _quat([0.92, 0.12, 0.18, 0.31])
    q2, q2bar = _quat([0.85, 0.14, 0.18, 0.48]), _quat([0.86, 0.13, 0.17, 0.47])
    p1, p1bar = np.array([1.0, -2.0, 0.5]), np.array([1.1, -2.1, 0.4])
    iq1 = b.add_quaternion(q1); ip1 = b.add_block(p1); iq2 = b.add_quaternion(q2); ip2 = b.add_block([2.0, -1.0, 0.0])
    A = [rng.normal(size=(rows, 3)) for _ in range(3)]
    bb = 0.1 * rng.normal(size=rows)
    b.add_marginal([iq1, ip1, iq2], A, bb, [q1bar, p1bar, q2bar])
    if with_chain:
        b.add_factor(cabi.FACTOR_RELATIVE_POSE, [ip1, iq1, ip2, iq2], np.concatenate([[1, 0.5, -0.5, 0.99, 0.05, 0.0, 0.1], np.eye(6).ravel()]))
        b.add_factor(cabi.FACTOR_ABSOLUTE_POSE, [ip2, iq2], np.concatenate([[2, -1, 0, *q2bar], 2.0 * np.eye(6).ravel()]))
    return b.build(), (q1, q1bar, p1, p1bar, q2, q2bar, A, bb)


def test_marginal_cost_function_is_the_reference_formula():
    """residual = b + sum A_i Minus(x_bar_i, x_i); ambient Jacobian = A_i * MinusJacobian(x_i) (marginal_cost_function.cpp:72-110,
    orientation_3d_stamped.h:130-142); in the tangent space that is A_i itself, whatever the distance to x_bar."""
    p, (q1, q1bar, p1, p1bar, q2, q2bar, A, bb) = _problem(with_chain=False)
    res, jac = orc.cost_function(p, cabi.FACTOR_MARGINAL, 0, sizes=[4, 3, 4])
    rows = len(bb)
    expect = bb + A[0] @ _minus(q1bar, q1) + A[1] @ (p1 - p1bar) + A[2] @ _minus(q2bar, q2)
    assert np.abs(res[:rows] - expect).max() < 1e-14 and np.abs(res[rows:]).max() == 0.0
    for q, Ai, J in ((q1, A[0], jac[0]), (q2, A[2], jac[2])):
        x0, x1, x2, x3 = 2 * q
        Jm = np.array([[-x1, x0, x3, -x2], [-x2, -x3, x0, x1], [-x3, x2, -x1, x0]])
        assert np.abs(J[:rows] - Ai @ Jm).max() < 1e-14
    assert np.abs(jac[1][:rows] - A[1]).max() == 0.0
    r2, Jt = orc.linearize(p, cabi.FACTOR_MARGINAL, 0)
    assert np.abs(Jt[:rows] - np.hstack(A)).max() < 1e-14 and np.abs(Jt[rows:]).max() == 0.0


def test_marginal_prior_alone_is_solved_in_one_gauss_newton_step_by_the_oracle():
    """A full-rank square prior on vectors is linear: the minimiser is x_bar - A^-1 b."""
    rng = np.random.default_rng(5)
    b = ProblemBuilder()
    i1 = b.add_block([0.3, -0.2, 0.9]); i2 = b.add_block([1.0, 2.0, 3.0])
    M = rng.normal(size=(6, 6)) + 3 * np.eye(6)
    bb = rng.normal(size=6); xbar = [np.array([0.1, 0.2, 0.3]), np.array([1.5, 1.0, 2.0])]
    b.add_marginal([i1, i2], [M[:, :3], M[:, 3:]], bb, xbar)
    so, its, x = orc.solve(b.build())
    assert np.abs(x - (np.concatenate(xbar) - np.linalg.solve(M, bb))).max() < 1e-6
    assert so.final_cost < 1e-12


@pytest.mark.gpu
def test_marginal_factor_on_the_gpu_matches_the_oracle():
    from beam_slam_b200.solver import Solver
    p, _ = _problem()
    s = Solver(); s.set_problem(p)
    cost, res, grad = s.evaluate()
    co, ro, go = orc.evaluate(p)
    assert abs(cost - co) <= 1e-12 * co and np.abs(res - ro).max() <= 1e-12 * max(1.0, np.abs(ro).max())
    assert np.abs(grad - go).max() <= 1e-11 * max(1.0, np.abs(go).max())
    summ, its = s.solve(); x = s.get_values()
    so, ito, xo = orc.solve(p)
    assert summ.termination_type == so.termination_type and len(its) == len(ito)
    for a, c in zip(its, ito):
        assert abs(a.cost - c.cost) <= 1e-9 * max(abs(c.cost), 1e-30) + 1e-22
    assert np.abs(x - xo).max() < 1e-8


@pytest.mark.gpu
def test_marginal_factor_shape_is_checked():
    from beam_slam_b200.solver import BsnlsError, Solver
    p, _ = _problem()
    g = p.groups[cabi.FACTOR_MARGINAL]
    g["consts"][0, 0] = 2          # says two blocks, lists three
    p = cabi.FlatProblem(p.block_size, p.block_manifold, p.block_constant, p.block_group, p.block_values, p.cameras, p.groups)
    with pytest.raises(BsnlsError):
        Solver().set_problem(p)


def test_marginalize_next_reproduces_the_reference_golden_values():
    """test_marginalize_variables.cpp:358-414 (MarginalizeNext): the literal terms and the literal expected marginal."""
    from beam_slam_b200 import host
    A1 = [[0.99999999999999922284, 4.4999993911720714834e-08, -2.9999995598828377297e-08],
          [-3.181980062078038074e-08, 0.70710678118654701763, 1.0606600528428877794e-08],
          [1.7320505793505525105e-08, -8.6602525498080673572e-09, 0.57735026918962550901]]
    b1 = [-2.9999995786018886696e-08, -4.2426400911723613558e-08, -5.1961516896187549911e-08]
    A21 = [[0.99999999999996114219, -1.8482708254510815671e-07, -2.873543621662033587e-07],
           [1.3069243487082160549e-07, 0.70710678118650927004, -2.5352115484711390536e-07],
           [1.6590414383954588118e-07, 2.0699913566568639567e-07, 0.57735026918958520792]]
    A22 = [[-0.91999992754510684367, -0.39191852892782985673, -2.8735440640859089001e-07],
           [0.27712824947756498073, -0.6505381874582485402, -2.5352112979770691226e-07],
           [7.1505019336171038447e-08, 2.5546009342625186633e-07, -0.57735026918958520792]]
    b2 = [7.1706607563166841896e-07, -4.0638046747479327072e-07, 2.1341989211309879704e-07]
    variables, A, b = host.marginalize_next([([1], [A1], b1), ([3, 1], [A21, A22], b2)])
    assert variables == [3]
    A_expected = np.array([[-0.686835139329528, 0.064384601986636, 0.000000153209328],
                           [0.000000000000000, -0.509885650799691, 0.000000079984512],
                           [0.000000000000000, 0.000000000000000, 0.408248290463911]])
    b_expected = np.array([-0.000000497197868, 0.000000315186479, 0.000000114168337])
    assert np.abs(A[0] - A_expected).max() < 1e-9 and np.abs(b - b_expected).max() < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("which", [0, 1, 2])
def test_marginalizing_a_pose_leaves_the_solution_and_covariance_of_the_others_unchanged(which):
    """The check of test_marginalize_variables.cpp:416-560 (optimise, marginalise, re-optimise: values and covariances of the
    remaining variables agree to 1e-5) on a pose graph: pose 0 (prior + two links), pose 1 (three links) or pose 2 (four
    links: the marginal has 18 rows and leaves as two constraints of 15 and 3 rows)."""
    from beam_slam_b200 import host
    g = host.HashGraph.pose_chain(6, seed=4)
    o = host.HashGraph.synthetic_options(3)
    s, its = g.optimize(o)
    assert s.is_solution_usable()
    p = g.flatten(); uu = g.block_uuids(); x = g.current_values()
    # the two blocks of the pose to remove: the ones of the `which`-th stamp (pose k has stamp 10 + k)
    from beam_slam_b200.host import load_library
    import ctypes as C
    lib = load_library()
    def stamped(kind, k):
        buf = C.create_string_buffer(16); lib.bsfuse_uuid_stamped(kind.encode(), 10 + k, 0, buf); return buf.raw
    gone = [stamped("fuse_variables::Position3DStamped", which), stamped("fuse_variables::Orientation3DStamped", which)]
    assert all(u in uu for u in gone)
    keep = [b for b in range(len(p.block_size)) if uu[b] not in gone]
    before = {uu[b]: x[p.offsets[b]:p.offsets[b] + p.block_size[b]].copy() for b in keep}
    cov_before = dict(zip([uu[b] for b in keep], g.covariance([(b, b) for b in keep])))
    removed_c, added_c, removed_v = g.marginalize(gone)
    assert removed_v == 2 and removed_c == (4 if which == 2 else 3) and added_c == (2 if which == 2 else 1)
    s2, _ = g.optimize(o)
    assert s2.is_solution_usable()
    p2 = g.flatten(); uu2 = g.block_uuids(); x2 = g.current_values()
    assert len(p2.block_size) == len(p.block_size) - 2 and cabi.FACTOR_MARGINAL in p2.groups
    cov_after = dict(zip(uu2, g.covariance([(b, b) for b in range(len(p2.block_size))])))
    for b in range(len(p2.block_size)):
        assert np.abs(x2[p2.offsets[b]:p2.offsets[b] + p2.block_size[b]] - before[uu2[b]]).max() < 1e-5
        assert np.abs(cov_after[uu2[b]] - cov_before[uu2[b]]).max() < 1e-5
    with pytest.raises(KeyError):
        g.marginalize(gone)          # already removed: std::out_of_range


@pytest.mark.gpu
def test_marginal_that_does_not_fit_the_gpu_factor_is_refused():
    """A keyframe of a visual window touches hundreds of landmarks: more blocks than BSNLS_MARGINAL_MAX_BLOCKS."""
    from beam_slam_b200 import host
    g = host.HashGraph.synthetic(1)
    p = g.flatten(); uu = g.block_uuids()
    pose_q = int(p.groups[cabi.FACTOR_REPROJECTION]["blocks"][0, 0])
    with pytest.raises(ValueError):
        g.marginalize([uu[pose_q]])


@pytest.mark.gpu
def test_device_linearisation_matches_the_oracle_for_every_factor_type():
    """bsnls_linearize against the oracle's restatement of detail::linearize (marginalize_variables.cpp:281-399) on a
    visual-inertial window (reprojection with Cauchy loss, IMU deltas, the 15-DoF prior) and on a refinement graph
    (relative poses with Cauchy loss, the pose prior)."""
    from beam_slam_b200 import host
    from beam_slam_b200.solver import Solver
    for cfg, scale in ((1, 1.0), (3, 0.02)):
        p = host.HashGraph.synthetic(cfg, scale=scale).flatten()
        s = Solver(host.HashGraph.synthetic_options(cfg)); s.set_problem(p)
        for t, g in p.groups.items():
            n = len(g["res_index"]); pick = sorted(set([0, n // 2, n - 1]))
            res, jac = s.linearize([int(g["res_index"][i]) for i in pick])
            nr = cabi.FACTOR_NRES[t]
            for q, i in enumerate(pick):
                ro, Jo = orc.linearize(p, t, i)
                assert np.abs(res[q, :nr] - ro).max() <= 1e-11 * max(1.0, np.abs(ro).max())
                col = 0
                for b, blk in enumerate(g["blocks"][i]):
                    if blk < 0 or p.block_constant[blk]:
                        continue
                    assert np.abs(jac[q, :nr, 3 * b:3 * b + 3] - Jo[:, col:col + 3]).max() <= 1e-9 * max(1.0, np.abs(Jo).max())
                    col += 3
                assert col == Jo.shape[1]


def test_gravity_alignment_and_absolute_orientation_cost_functions():
    """GravityAlignmentCostFunctor (bs_constraints/global/gravity_alignment_cost_functor.h:53-66): A [R(q) g]_xy;
    NormalPriorOrientation3DCostFunctor (normal_prior_orientation_3d_cost_functor.h:92-125): A AngleAxis(mean^-1 * q)."""
    q = _quat([0.9, 0.2, -0.3, 0.1]); g = np.array([0.1, -0.2, -9.7]); A2 = np.array([[2.0, 0.5], [0.0, 3.0]])
    mean = _quat([0.95, 0.1, -0.2, 0.05]); A3 = np.triu(np.arange(1.0, 10.0).reshape(3, 3))
    b = ProblemBuilder(); iq = b.add_quaternion(q)
    b.add_factor(cabi.FACTOR_GRAVITY_ALIGNMENT, [iq], np.concatenate([g, A2.ravel()]))
    b.add_factor(cabi.FACTOR_ABSOLUTE_ORIENTATION, [iq], np.concatenate([mean, A3.ravel()]))
    p = b.build()
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    res, _ = orc.cost_function(p, cabi.FACTOR_GRAVITY_ALIGNMENT, 0)
    assert np.abs(res - A2 @ (R @ g)[:2]).max() < 1e-13
    res, _ = orc.cost_function(p, cabi.FACTOR_ABSOLUTE_ORIENTATION, 0)
    assert np.abs(res - A3 @ _minus(mean, q)).max() < 1e-13


@pytest.mark.gpu
def test_gravity_alignment_and_absolute_orientation_on_the_gpu_match_the_oracle():
    from beam_slam_b200.solver import Solver
    b = ProblemBuilder()
    iq = b.add_quaternion(_quat([0.9, 0.2, -0.3, 0.1])); iq2 = b.add_quaternion(_quat([0.8, -0.1, 0.3, 0.4]))
    b.add_factor(cabi.FACTOR_GRAVITY_ALIGNMENT, [iq], [0.0, 0.0, -9.81, 1.0, 0.2, 0.0, 1.5])
    b.add_factor(cabi.FACTOR_GRAVITY_ALIGNMENT, [iq2], [0.1, 0.0, -9.8, 1.0, 0.0, 0.0, 1.0], loss_type=cabi.LOSS_CAUCHY, loss_param=1.0)
    b.add_factor(cabi.FACTOR_ABSOLUTE_ORIENTATION, [iq], np.concatenate([_quat([0.95, 0.1, -0.2, 0.05]), np.eye(3).ravel()]))
    b.add_factor(cabi.FACTOR_RELATIVE_ORIENTATION, [iq, iq2], np.concatenate([_quat([0.98, 0.0, 0.1, 0.15]), 2 * np.eye(3).ravel()]))
    p = b.build()
    s = Solver(); s.set_problem(p)
    cost, res, grad = s.evaluate(); co, ro, go = orc.evaluate(p)
    assert abs(cost - co) <= 1e-12 * co and np.abs(res - ro).max() <= 1e-12 * max(1.0, np.abs(ro).max()) and np.abs(grad - go).max() <= 1e-10 * max(1.0, np.abs(go).max())
    summ, its = s.solve(); so, ito, xo = orc.solve(p)
    assert len(its) == len(ito) and np.abs(s.get_values() - xo).max() < 1e-8
    for a, c in zip(its, ito):
        assert abs(a.cost - c.cost) <= 1e-9 * max(abs(c.cost), 1e-30) + 1e-22


@pytest.mark.gpu
def test_reference_marginalize_variables_scenario():
    """test_marginalize_variables.cpp:416-560 with its literal graph: optimise, marginalise x1 (the prior and x1-x2 go, one
    marginal constraint comes), re-optimise; x2, x3, l1 and their covariances agree to 1e-5 with the values before."""
    import ctypes as C
    from beam_slam_b200 import host
    g = host.HashGraph.orientation_chain()
    o = cabi.Options.defaults()
    s, _ = g.optimize(o)
    assert s.is_solution_usable()
    p = g.flatten(); uu = g.block_uuids(); x = g.current_values()
    buf = C.create_string_buffer(16); host.load_library().bsfuse_uuid_stamped(b"fuse_variables::Orientation3DStamped", 1, 0, buf)
    x1 = buf.raw
    keep = [b for b in range(4) if uu[b] != x1]
    assert len(keep) == 3
    before = {uu[b]: x[p.offsets[b]:p.offsets[b] + 4].copy() for b in keep}
    cov_before = dict(zip([uu[b] for b in keep], g.covariance([(b, b) for b in keep])))
    assert g.marginalize([x1]) == (2, 1, 1)
    s2, _ = g.optimize(o)
    assert s2.is_solution_usable()
    p2 = g.flatten(); uu2 = g.block_uuids(); x2 = g.current_values()
    assert len(p2.block_size) == 3
    cov_after = g.covariance([(b, b) for b in range(3)])
    for b in range(3):
        assert np.abs(x2[p2.offsets[b]:p2.offsets[b] + 4] - before[uu2[b]]).max() < 1e-5
        assert np.abs(cov_after[b] - cov_before[uu2[b]]).max() < 1e-5


@pytest.mark.gpu
def test_successive_marginalisation_relinearises_the_marginal_prior():
    """Marginalise pose 0, then pose 1: the second step linearises the MarginalConstraint the first one left (it is tied to
    pose 1), as marginalize_variables.cpp:203-224 does with any connected constraint. The optimum and the covariances of the
    remaining poses still agree with the full graph's (1e-5); marginalising both poses in one call gives the same."""
    import ctypes as C
    from beam_slam_b200 import host
    lib = host.load_library()
    def stamped(kind, k):
        buf = C.create_string_buffer(16); lib.bsfuse_uuid_stamped(kind.encode(), 10 + k, 0, buf); return buf.raw
    def pose(k):
        return [stamped("fuse_variables::Position3DStamped", k), stamped("fuse_variables::Orientation3DStamped", k)]
    o = host.HashGraph.synthetic_options(3)
    results = []
    for steps in ([pose(0), pose(1)], [pose(0) + pose(1)]):
        g = host.HashGraph.pose_chain(7, seed=9)
        assert g.optimize(o)[0].is_solution_usable()
        p = g.flatten(); uu = g.block_uuids(); x = g.current_values()
        full = {uu[b]: x[p.offsets[b]:p.offsets[b] + p.block_size[b]].copy() for b in range(len(uu))}
        cov_full = dict(zip(uu, g.covariance([(b, b) for b in range(len(uu))])))
        for gone in steps:
            g.marginalize(gone)
            assert g.optimize(o)[0].is_solution_usable()
        p2 = g.flatten(); uu2 = g.block_uuids(); x2 = g.current_values()
        assert len(uu2) == len(uu) - 4 and cabi.FACTOR_MARGINAL in p2.groups
        cov2 = g.covariance([(b, b) for b in range(len(uu2))])
        for b in range(len(uu2)):
            assert np.abs(x2[p2.offsets[b]:p2.offsets[b] + p2.block_size[b]] - full[uu2[b]]).max() < 1e-5
            assert np.abs(cov2[b] - cov_full[uu2[b]]).max() < 1e-5
        results.append({uu2[b]: x2[p2.offsets[b]:p2.offsets[b] + p2.block_size[b]].copy() for b in range(len(uu2))})
    for u, v in results[0].items():
        assert np.abs(v - results[1][u]).max() < 1e-6


def test_snapshot_keeps_the_new_factor_types(tmp_path):
    """The marginal prior (ten block slots, -1 padded), gravity alignment and the orientation prior survive
    bsnls_problem_save / _load and the numpy reader bit for bit, and the oracle gives the same cost on the copy."""
    p, _ = _problem()
    b = ProblemBuilder()
    iq = b.add_quaternion(_quat([0.9, 0.2, -0.3, 0.1]))
    b.add_factor(cabi.FACTOR_GRAVITY_ALIGNMENT, [iq], [0.0, 0.0, -9.81, 1.0, 0.2, 0.0, 1.5])
    b.add_factor(cabi.FACTOR_ABSOLUTE_ORIENTATION, [iq], np.concatenate([_quat([0.95, 0.1, -0.2, 0.05]), np.eye(3).ravel()]))
    for k, prob in enumerate((p, b.build())):
        path = tmp_path / f"p{k}.bsnls"
        prob.save(path)
        for q in (cabi.FlatProblem.load(path), cabi.FlatProblem.load_numpy(path)):
            assert sorted(q.groups) == sorted(prob.groups)
            for t in prob.groups:
                assert np.array_equal(q.groups[t]["blocks"], prob.groups[t]["blocks"]) and np.array_equal(q.groups[t]["consts"], prob.groups[t]["consts"])
            assert orc.evaluate(q)[0] == orc.evaluate(prob)[0]
