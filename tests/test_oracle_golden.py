"""Pins the CPU oracle (oracle/) to the golden vectors and known-answer tests the reference holds for the path
(SURVEY.md section 8c). CPU only."""
import numpy as np
import pytest

import oracle_lib as orc
from beam_slam_b200 import cabi
from beam_slam_b200.builder import ProblemBuilder


def test_linearize_golden_relative_orientation():
    """dependencies/fuse/fuse_constraints/test/test_marginalize_variables.cpp:298-356 — A0, A1, b at 1e-9."""
    b = ProblemBuilder()
    x1 = b.add_quaternion([0.927362, 0.1, 0.2, 0.3])
    x2 = b.add_quaternion([0.848625, 0.13798, 0.175959, 0.479411])
    cov = np.diag([1.0, 2.0, 3.0])
    A = np.linalg.cholesky(np.linalg.inv(cov)).T  # cov.inverse().llt().matrixU()
    b.add_factor(cabi.FACTOR_RELATIVE_ORIENTATION, [x1, x2], np.concatenate([[0.979795897, 0.0, 0.0, 0.2], A.ravel()]))
    res, J = orc.linearize(b.build(), cabi.FACTOR_RELATIVE_ORIENTATION, 0)
    A0 = np.array([[-0.91999992754510684367, -0.39191852892782985673, -2.8735440640859089001e-07],
                   [0.27712824947756498073, -0.65053818745824854020, -2.5352112979770691226e-07],
                   [7.1505019336171038447e-08, 2.5546009342625186633e-07, -0.57735026918958520792]])
    A1 = np.array([[0.99999999999996114219, -1.8482708254510815671e-07, -2.873543621662033587e-07],
                   [1.3069243487082160549e-07, 0.70710678118650927004, -2.5352115484711390536e-07],
                   [1.6590414383954588118e-07, 2.0699913566568639567e-07, 0.57735026918958520792]])
    bb = np.array([7.1706607563166841896e-07, -4.0638046747479327072e-07, 2.1341989211309879704e-07])
    assert np.abs(J[:, 0:3] - A0).max() < 1e-9
    assert np.abs(J[:, 3:6] - A1).max() < 1e-9
    assert np.abs(res - bb).max() < 1e-9


def test_quaternion_plus_minus_golden():
    """dependencies/fuse/fuse_variables/test/test_orientation_3d_stamped.cpp:142-175 (1e-5)."""
    x = [0.842614977, 0.2, 0.3, 0.4]
    d = [0.15, -0.2, 0.433012702]
    r = orc.quat_plus(x, d)
    assert np.abs(r - [0.745561, 0.360184, 0.194124, 0.526043]).max() < 1e-5
    m = orc.quat_minus(x, [0.745561, 0.360184, 0.194124, 0.526043])
    assert np.abs(m - d).max() < 1e-5


def _rand_quat(rng):
    q = rng.normal(size=4)
    return q / np.linalg.norm(q)


def _quat_to_R(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _reproj_problem(q, t, P, K, T_cb, pixel, w=1.0, loss=cabi.LOSS_NONE, a=0.0):
    b = ProblemBuilder()
    iq = b.add_quaternion(q)
    it = b.add_block(t)
    ip = b.add_landmark(P)
    cam = b.add_camera(K[0], K[1], K[2], K[3], T_cb, w, loss, a)
    b.add_factor(cabi.FACTOR_REPROJECTION, [iq, it, ip], pixel, aux=cam)
    return b.build()


def test_reprojection_zero_residual_at_constructed_point():
    """bs_models/tests/reprojection_test.cpp:13-74 with bs_models/tests/data/intrinsics.json: a point built by
    back-projecting a pixel 10 m along its ray reprojects onto it (pinhole part of the model)."""
    K = (458.654, 457.296, 367.215, 248.375)
    T_imu_cam = np.array([[0.0148655429818, -0.999880929698, 0.00414029679422, -0.0216401454975],
                          [0.999557249008, 0.0149672133247, 0.025715529948, -0.064676986768],
                          [-0.0257744366974, 0.00375618835797, 0.999660727178, 0.00981073058949],
                          [0, 0, 0, 1.0]])
    T_world_cam = np.eye(4)
    T_world_cam[:3, 3] = [5, 5, 5]
    T_world_imu = T_world_cam @ np.linalg.inv(T_imu_cam)
    R = T_world_imu[:3, :3]
    # rotation matrix -> quaternion
    w = np.sqrt(max(0.0, 1 + np.trace(R))) / 2
    q = np.array([w, (R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w)])
    pixel = np.array([240.0, 376.0])
    ray = np.array([(pixel[0] - K[2]) / K[0], (pixel[1] - K[3]) / K[1], 1.0])
    point_world = (T_world_cam @ np.append(10 * ray / np.linalg.norm(ray), 1.0))[:3]
    p = _reproj_problem(q, T_world_imu[:3, 3], point_world, K, np.linalg.inv(T_imu_cam), pixel)
    res, _ = orc.cost_function(p, cabi.FACTOR_REPROJECTION, 0)
    assert np.abs(res).max() < 1e-3  # the reference's tolerance; R is only orthonormal to ~1e-12
    assert np.abs(res).max() < 1e-6


def test_reprojection_fd_jacobian_matches_numeric_tangent_jacobian():
    """bs_constraints/tests/euclidean_reprojection_test.cpp:31-197: the hand-rolled (finite-difference) Jacobian
    agrees with an exact one to 1e-5 in tangent space, same structure (2 x 9)."""
    rng = np.random.default_rng(7)
    for _ in range(50):
        q = _rand_quat(rng)
        t = rng.uniform(1, 10, 3)
        K = (rng.uniform(10, 100),) * 2 + tuple(rng.uniform(100, 1000, 2))
        Rcb = _quat_to_R(_rand_quat(rng))
        T_cb = np.eye(4)
        T_cb[:3, :3] = Rcb
        T_cb[:3, 3] = rng.uniform(0, 1, 3)
        P_cam = rng.uniform(5, 10) * (lambda v: v / np.linalg.norm(v))(rng.uniform(0.1, 1, 3))
        T_wb = np.eye(4)
        T_wb[:3, :3] = _quat_to_R(q)
        T_wb[:3, 3] = t
        P_w = (T_wb @ np.linalg.inv(T_cb) @ np.append(P_cam, 1))[:3]
        pix = np.array([K[0] * P_cam[0] / P_cam[2] + K[2], K[1] * P_cam[1] / P_cam[2] + K[3]]) + 10
        prob = _reproj_problem(q, t, P_w, K, T_cb, pix)
        res, J = orc.linearize(prob, cabi.FACTOR_REPROJECTION, 0)
        assert J.shape == (2, 9)
        assert np.allclose(res, [10, 10], atol=1e-6)

        def f(dq, dt, dP):
            pr = _reproj_problem(orc.quat_plus(q, dq), t + dt, P_w + dP, K, T_cb, pix)
            return orc.cost_function(pr, cabi.FACTOR_REPROJECTION, 0)[0]
        Jn = np.zeros((2, 9))
        h = 1e-6
        for k in range(9):
            d = np.zeros(9)
            d[k] = h
            Jn[:, k] = (f(d[:3], d[3:6], d[6:]) - f(-d[:3], -d[3:6], -d[6:])) / (2 * h)
        assert np.abs(J - Jn).max() < 1e-5 * max(1.0, np.abs(Jn).max())


COV15 = None


def _cov15():
    c = np.zeros((15, 15))
    for i in range(15):
        for j in range(15):
            if i == j:
                c[i, j] = i + 1.0
            elif i == 0 or j == 0:
                c[i, j] = 0.1 * max(i, j)
            else:
                c[i, j] = 1.5 - 0.1 * (abs(i - j) - 1)
    return c


def test_cov15_literal():
    """first rows of the literal in bs_constraints/tests/absolute_imu_state_3d_stamped_constraint_test.cpp:36-52"""
    c = _cov15()
    assert np.allclose(c[0, :5], [1.0, 0.1, 0.2, 0.3, 0.4])
    assert np.allclose(c[1, :6], [0.1, 2.0, 1.5, 1.4, 1.3, 1.2])
    assert np.allclose(c[14, 10:], [1.2, 1.3, 1.4, 1.5, 15.0])
    assert np.allclose(c, c.T)


def _tangent_information(problem, factors):
    """sum J^T J over all factors in the order blocks were added (ceres::Covariance's normal matrix)."""
    nb = len(problem.block_size)
    off = np.concatenate([[0], np.cumsum([3] * nb)])
    H = np.zeros((off[-1], off[-1]))
    for t, i in factors:
        _, J = orc.linearize(problem, t, i)
        blocks = problem.groups[t]["blocks"][i]
        cols = np.concatenate([np.arange(off[b], off[b + 1]) for b in blocks])
        H[np.ix_(cols, cols)] += J.T @ J
    return H


def test_absolute_imu_state_prior_optimization():
    """bs_constraints/tests/absolute_imu_state_3d_stamped_constraint_test.cpp:85-298: one 15-DoF prior, solve ->
    mean (1e-3 quaternion, 1e-5 rest) and the covariance comes back (1e-5)."""
    cov = _cov15()
    mean = np.array([1.0, 0, 0, 0, 1, 2, 3, 1, 2, 3, 0.1, 0.2, 0.3, 0.1, 0.2, 0.3])
    A = np.linalg.cholesky(np.linalg.inv(cov)).T
    b = ProblemBuilder()
    ids = [b.add_quaternion([0.952, 0.038, -0.189, 0.239]), b.add_block([1.5, -3.0, 10.0]),
           b.add_block([1.5, -3.0, 10.0]), b.add_block([0.15, -0.30, 1.0]), b.add_block([0.15, -0.30, 1.0])]
    b.add_factor(cabi.FACTOR_IMU_PRIOR, ids, np.concatenate([mean, A.ravel()]))
    p = b.build()
    s, its, x = orc.solve(p)
    assert s.termination_type == cabi.CONVERGENCE
    assert np.abs(x[:4] - mean[:4]).max() < 1e-3
    assert np.abs(x[4:] - mean[4:]).max() < 1e-5
    H = _tangent_information(p.with_values(x), [(cabi.FACTOR_IMU_PRIOR, 0)])
    assert np.abs(np.linalg.inv(H) - cov).max() < 1e-5


def test_relative_pose_3d_optimization_and_covariance():
    """dependencies/fuse/fuse_constraints/test/test_relative_pose_3d_stamped_constraint.cpp:124-306: prior at the
    origin + 1 m relative pose; optimum and both marginal covariances (1e-9)."""
    b = ProblemBuilder()
    o1 = b.add_quaternion([0.952, 0.038, -0.189, 0.239])
    p1 = b.add_block([1.5, -3.0, 10.0])
    o2 = b.add_quaternion([0.944, -0.128, 0.145, -0.269])
    p2 = b.add_block([-1.5, 3.0, -10.0])
    I6 = np.eye(6).ravel()
    b.add_factor(cabi.FACTOR_ABSOLUTE_POSE, [p1, o1], np.concatenate([[0, 0, 0, 1, 0, 0, 0], I6]))
    b.add_factor(cabi.FACTOR_RELATIVE_POSE, [p1, o1, p2, o2], np.concatenate([[1, 0, 0, 1, 0, 0, 0], I6]))
    p = b.build()
    s, its, x = orc.solve(p)
    assert s.is_solution_usable()
    q1, t1, q2, t2 = x[0:4], x[4:7], x[7:11], x[11:14]
    assert np.abs(t1).max() < 1e-5 and np.abs(t2 - [1, 0, 0]).max() < 1e-5
    assert np.abs(q1 - [1, 0, 0, 0]).max() < 1e-3 and np.abs(q2 - [1, 0, 0, 0]).max() < 1e-3
    # exact optimum for the covariance check, as ceres::Covariance evaluates at the converged point
    exact = p.with_values([1, 0, 0, 0, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0])
    H = _tangent_information(exact, [(cabi.FACTOR_ABSOLUTE_POSE, 0), (cabi.FACTOR_RELATIVE_POSE, 0)])
    C = np.linalg.inv(H)  # block order: o1, p1, o2, p2
    pose1 = np.ix_([3, 4, 5, 0, 1, 2], [3, 4, 5, 0, 1, 2])
    pose2 = np.ix_([9, 10, 11, 6, 7, 8], [9, 10, 11, 6, 7, 8])
    assert np.abs(C[pose1] - np.eye(6)).max() < 1e-9
    expected2 = np.array([[2, 0, 0, 0, 0, 0], [0, 3, 0, 0, 0, 1], [0, 0, 3, 0, -1, 0],
                          [0, 0, 0, 2, 0, 0], [0, 0, -1, 0, 2, 0], [0, 1, 0, 0, 0, 2.0]])
    assert np.abs(C[pose2] - expected2).max() < 1e-9


def test_hash_graph_optimize_known_answer_and_hold():
    """dependencies/fuse/fuse_graphs/test/test_hash_graph.cpp:659-737 uses scalar example variables; the same two
    checks on the 3-vector prior: optimize() reaches the prior means; a held variable does not move."""
    b = ProblemBuilder()
    v1 = b.add_block([1.0, 1.0, 1.0])
    v2 = b.add_block([2.0, 2.0, 2.0], constant=True)
    I3 = np.eye(3).ravel()
    b.add_factor(cabi.FACTOR_ABSOLUTE_VEC3, [v1], np.concatenate([[5.0, -3.0, 0.5], I3]))
    b.add_factor(cabi.FACTOR_ABSOLUTE_VEC3, [v2], np.concatenate([[7.0, 7.0, 7.0], I3]))
    s, its, x = orc.solve(b.build())
    assert np.abs(x[:3] - [5.0, -3.0, 0.5]).max() < 1e-7
    assert np.array_equal(x[3:], [2.0, 2.0, 2.0])
    assert abs(s.fixed_cost - 0.5 * 75.0) < 1e-12


def test_cauchy_corrector_against_fuse_linearize_formula():
    """dependencies/fuse/fuse_constraints/src/marginalize_variables.cpp:361-396 applied to a reprojection block
    with CauchyLoss(5): r, J scaled by sqrt(rho')."""
    rng = np.random.default_rng(3)
    q = _rand_quat(rng)
    K = (259.196, 258.951, 408.532, 316.260)
    args = (q, [0.1, 0.2, 0.3], [1.0, 2.0, 9.0], K, np.eye(4), [400.0, 300.0])
    r0, J0 = orc.linearize(_reproj_problem(*args), cabi.FACTOR_REPROJECTION, 0)
    r1, J1 = orc.linearize(_reproj_problem(*args, 1.0, cabi.LOSS_CAUCHY, 5.0), cabi.FACTOR_REPROJECTION, 0)
    s = r0 @ r0
    rho1 = 1.0 / (1.0 + s / 25.0)
    assert np.allclose(r1, np.sqrt(rho1) * r0, rtol=1e-14)
    assert np.allclose(J1, np.sqrt(rho1) * J0, rtol=1e-14)
    cost, _, _ = orc.evaluate(_reproj_problem(*args, 1.0, cabi.LOSS_CAUCHY, 5.0))
    assert abs(cost - 0.5 * 25.0 * np.log(1 + s / 25.0)) < 1e-12 * max(1, cost)


def simple_two_state_problem():
    """bs_models/tests/imu_preintegration_tests.cpp:292-404 (Simple2StateFG): two IMU states (q, p, v, bg, ba) with literal
    values; priors at the origin on state 1 (pose, velocity, both biases; unit covariance) and relative constraints from
    state 1 to state 2 (pose +1 m in x, velocity +1 m/s in x, biases +0.001 in x; unit covariance). Returns the flat
    problem and the optimum the reference asserts, block by block."""
    b = ProblemBuilder()
    o1 = b.add_quaternion([0.952, 0.038, -0.189, 0.239]); p1 = b.add_block([1.5, -3.0, 1.0]); v1 = b.add_block([1.5, -3.0, 1.0])
    bg1 = b.add_block([4e-5, 5e-5, 6e-5]); ba1 = b.add_block([1e-5, 2e-5, 3e-5])
    o2 = b.add_quaternion([0.944, -0.128, 0.145, -0.269]); p2 = b.add_block([-1.5, 3.0, -1.0]); v2 = b.add_block([-1.5, 3.0, -1.0])
    bg2 = b.add_block([4e-5, 5e-5, 6e-5]); ba2 = b.add_block([1e-5, 2e-5, 3e-5])
    I6, I3 = np.eye(6).ravel(), np.eye(3).ravel()
    b.add_factor(cabi.FACTOR_ABSOLUTE_POSE, [p1, o1], np.concatenate([[0, 0, 0, 1, 0, 0, 0], I6]))
    for blk in (v1, bg1, ba1):
        b.add_factor(cabi.FACTOR_ABSOLUTE_VEC3, [blk], np.concatenate([[0, 0, 0], I3]))
    b.add_factor(cabi.FACTOR_RELATIVE_POSE, [p1, o1, p2, o2], np.concatenate([[1, 0, 0, 1, 0, 0, 0], I6]))
    b.add_factor(cabi.FACTOR_RELATIVE_VEC3, [v1, v2], np.concatenate([[1, 0, 0], I3]))
    for x1, x2 in ((bg1, bg2), (ba1, ba2)):
        b.add_factor(cabi.FACTOR_RELATIVE_VEC3, [x1, x2], np.concatenate([[0.001, 0, 0], I3]))
    expect = [[1, 0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0],
              [1, 0, 0, 0], [1, 0, 0], [1, 0, 0], [0.001, 0, 0], [0.001, 0, 0]]
    return b.build(), expect


def check_two_state_optimum(x, expect):
    """Tolerances of imu_preintegration_tests.cpp:418-466: 1e-3 on quaternion components, 1e-5 on everything else."""
    o = 0
    for e in expect:
        got = np.asarray(x[o:o + len(e)])
        o += len(e)
        assert np.abs(got - e).max() < (1e-3 if len(e) == 4 else 1e-5), (got, e)


def test_simple_two_state_graph_reaches_the_reference_optimum():
    p, expect = simple_two_state_problem()
    s, its, x = orc.solve(p)
    assert s.termination_type == cabi.CONVERGENCE
    check_two_state_optimum(x, expect)


def test_sparse_and_dense_reduced_system_paths_agree():
    """The block-sparse reduced system + sparse Cholesky (oracle/oracle_sparse.h, used for configs 4 and 5) against the
    dense path on graphs small enough for both: same blocks, same structural pattern, same iterations."""
    from beam_slam_b200 import host
    for cfg, scale in ((2, 0.5), (4, 0.05)):
        g = host.HashGraph.synthetic(cfg, scale=scale)
        p, o = g.flatten(), host.HashGraph.synthetic_options(cfg)
        try:
            orc.set_linear_mode(orc.LINEAR_DENSE)
            sd, itd, xd = orc.solve(p, o)
            orc.set_linear_mode(orc.LINEAR_SPARSE)
            ss, its, xs = orc.solve(p, o)
        finally:
            orc.set_linear_mode(orc.LINEAR_AUTO)
        assert sd.message == ss.message and len(itd) == len(its)
        for a, b in zip(itd, its):
            assert abs(a.cost - b.cost) <= 1e-9 * b.cost
        assert abs(itd[1].step_norm - its[1].step_norm) <= 1e-9 * its[1].step_norm
        S, rhs = orc.reduced_system(p, 1e4, o)
        r, c, v, rh = orc.reduced_blocks(p, 1e4, o)
        nf = S.shape[0] // 3
        Sb = S.reshape(nf, 3, nf, 3).transpose(0, 2, 1, 3)
        assert np.abs(Sb[r, c] - v).max() <= 1e-12 * np.abs(S).max()
        mask = np.zeros((nf, nf), bool)
        mask[r, c] = True
        mask[c, r] = True
        assert np.abs(Sb[~mask]).max() == 0.0          # nothing outside the structural pattern
        assert np.abs(rh - rhs).max() <= 1e-12 * np.abs(rhs).max()


ONE_PARAMETER_LOSSES = [cabi.LOSS_CAUCHY, cabi.LOSS_HUBER, cabi.LOSS_SOFTLONE, cabi.LOSS_ARCTAN, cabi.LOSS_TUKEY, cabi.LOSS_DCS,
                        cabi.LOSS_FAIR, cabi.LOSS_GEMAN_MCCLURE, cabi.LOSS_WELSCH]


def test_loss_functions_like_the_reference_tests():
    """dependencies/fuse/fuse_loss/test/test_loss_function.cpp: (i) :63-103 rho' and rho'' agree with symmetric finite
    differences of rho (h = 1e-4, 1e-6) and rho' >= 0; (ii) :112-168 the values at s = 0 of the four losses fuse defines
    itself (DCS [0, 1, 0]; Fair [0, 1, -inf]; Geman-McClure [0, 1, -2/b]; Welsch [0, 1, -1/b], a = 0.7)."""
    h = 1e-4
    for lt in ONE_PARAMETER_LOSSES:
        for s in (0.357, 1.792, 7.351, 33.0):          # both sides of every threshold for a = 0.7 and a = 2.5
            for a in (0.7, 2.5):
                if lt in (cabi.LOSS_HUBER, cabi.LOSS_TUKEY) and abs(s - a * a) < 2 * h:
                    continue
                if lt == cabi.LOSS_DCS and abs(s - a) < 2 * h:
                    continue
                rho, fwd, bwd = orc.loss(lt, a, s), orc.loss(lt, a, s + h), orc.loss(lt, a, s - h)
                assert rho[1] >= 0.0
                assert abs((fwd[0] - bwd[0]) / (2 * h) - rho[1]) < 1e-6, (lt, a, s)
                assert abs((fwd[0] - 2 * rho[0] + bwd[0]) / (h * h) - rho[2]) < 1e-6, (lt, a, s)
    a = 0.7
    assert np.abs(orc.loss(cabi.LOSS_DCS, a, 0.0) - [0, 1, 0]).max() < 1e-6
    r = orc.loss(cabi.LOSS_FAIR, a, 0.0)
    assert abs(r[0]) < 1e-6 and abs(r[1] - 1) < 1e-6 and r[2] < -1e300
    assert np.abs(orc.loss(cabi.LOSS_GEMAN_MCCLURE, a, 0.0) - [0, 1, -2 / (a * a)]).max() < 1e-6
    assert np.abs(orc.loss(cabi.LOSS_WELSCH, a, 0.0) - [0, 1, -1 / (a * a)]).max() < 1e-6
