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
