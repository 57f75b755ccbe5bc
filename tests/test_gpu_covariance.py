"""HashGraph::getCovariance on the GPU (bsnls_get_covariance): blocks of the inverse normal matrix at the current values.

  * the reference's own expectation: dependencies/fuse/fuse_constraints/test/test_relative_pose_3d_stamped_constraint.cpp
    :233-306 — prior at the origin + 1 m relative pose: Cov(pose1) = I, Cov(pose2) = the literal 6x6, both at 1e-9
    (tangent space, fuse's default); and the 15-DoF prior of
    bs_constraints/tests/absolute_imu_state_3d_stamped_constraint_test.cpp:231-298: the covariance that built the prior
    comes back (1e-5);
  * a whole VIO window (landmarks eliminated, IMU factors, prior): pose / velocity / bias blocks against the dense inverse of
    J^T J assembled from the oracle's per-factor linearisation."""
import numpy as np
import pytest

import oracle_lib as orc
import test_oracle_golden as tg
from beam_slam_b200 import cabi, host
from beam_slam_b200.builder import ProblemBuilder
from beam_slam_b200.solver import BsnlsError, Solver

pytestmark = pytest.mark.gpu


def _cov_matrix(s, blocks):
    """Joint tangent covariance of `blocks` assembled from pairwise requests."""
    pairs = [(a, b) for a in blocks for b in blocks]
    out = s.covariance(pairs)
    n = len(blocks)
    C = np.zeros((3 * n, 3 * n))
    for (i, a), (j, b) in [((i, a), (j, b)) for i, a in enumerate(blocks) for j, b in enumerate(blocks)]:
        C[3 * i:3 * i + 3, 3 * j:3 * j + 3] = out[i * n + j]
    return C


def test_relative_pose_covariances_match_the_reference_literals():
    b = ProblemBuilder()
    o1 = b.add_quaternion([1.0, 0, 0, 0]); p1 = b.add_block([0.0, 0, 0])      # the optimum the reference reaches (:221-231)
    o2 = b.add_quaternion([1.0, 0, 0, 0]); p2 = b.add_block([1.0, 0, 0])
    I6 = np.eye(6).ravel()
    b.add_factor(cabi.FACTOR_ABSOLUTE_POSE, [p1, o1], np.concatenate([[0, 0, 0, 1, 0, 0, 0], I6]))
    b.add_factor(cabi.FACTOR_RELATIVE_POSE, [p1, o1, p2, o2], np.concatenate([[1, 0, 0, 1, 0, 0, 0], I6]))
    s = Solver()
    s.set_problem(b.build())
    assert np.abs(_cov_matrix(s, [p1, o1]) - np.eye(6)).max() < 1e-9
    expected2 = np.array([[2, 0, 0, 0, 0, 0], [0, 3, 0, 0, 0, 1], [0, 0, 3, 0, -1, 0],
                          [0, 0, 0, 2, 0, 0], [0, 0, -1, 0, 2, 0], [0, 1, 0, 0, 0, 2.0]])
    assert np.abs(_cov_matrix(s, [p2, o2]) - expected2).max() < 1e-9
    # ambient space: the quaternion block is lifted through the plus-Jacobian (4x4, rank 3)
    amb = s.covariance([(o2, o2)], tangent_space=False, sizes={o2: 4})[0]
    Jp = 0.5 * np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.0]])
    assert np.abs(amb - Jp @ expected2[3:, 3:] @ Jp.T).max() < 1e-9
    # after a solve from the reference's start values the covariance is the same to the solver's accuracy
    b2 = ProblemBuilder()
    o1 = b2.add_quaternion([0.952, 0.038, -0.189, 0.239]); p1 = b2.add_block([1.5, -3.0, 10.0])
    o2 = b2.add_quaternion([0.944, -0.128, 0.145, -0.269]); p2 = b2.add_block([-1.5, 3.0, -10.0])
    b2.add_factor(cabi.FACTOR_ABSOLUTE_POSE, [p1, o1], np.concatenate([[0, 0, 0, 1, 0, 0, 0], I6]))
    b2.add_factor(cabi.FACTOR_RELATIVE_POSE, [p1, o1, p2, o2], np.concatenate([[1, 0, 0, 1, 0, 0, 0], I6]))
    s2 = Solver()
    s2.set_problem(b2.build())
    s2.solve()
    assert np.abs(_cov_matrix(s2, [p2, o2]) - expected2).max() < 1e-6


def test_imu_prior_covariance_comes_back():
    cov = tg._cov15()
    mean = np.array([1.0, 0, 0, 0, 1, 2, 3, 1, 2, 3, 0.1, 0.2, 0.3, 0.1, 0.2, 0.3])
    A = np.linalg.cholesky(np.linalg.inv(cov)).T
    b = ProblemBuilder()
    ids = [b.add_quaternion(mean[:4]), b.add_block(mean[4:7]), b.add_block(mean[7:10]), b.add_block(mean[10:13]), b.add_block(mean[13:16])]
    b.add_factor(cabi.FACTOR_IMU_PRIOR, ids, np.concatenate([mean, A.ravel()]))
    s = Solver()
    s.set_problem(b.build())
    assert np.abs(_cov_matrix(s, ids) - cov).max() < 1e-5 * np.abs(cov).max()


def test_window_covariance_matches_dense_inverse_and_refuses_landmarks():
    g = host.HashGraph.synthetic(1, scale=0.2)            # 10 keyframes, 60 landmarks, IMU factors, prior
    p = g.flatten()
    s = Solver(host.HashGraph.synthetic_options(1))
    s.set_problem(p)
    s.solve()
    x = s.get_values()
    q = p.with_values(x)
    # dense J^T J over every non-constant block from the oracle's per-factor linearisation (tangent, loss-corrected)
    active = [b for b in range(len(p.block_size)) if not p.block_constant[b]]
    pos = {b: k for k, b in enumerate(active)}
    H = np.zeros((3 * len(active), 3 * len(active)))
    for t, grp in q.groups.items():
        for i in range(len(grp["res_index"])):
            _, J = orc.linearize(q, t, i)
            blocks = [int(b) for b in grp["blocks"][i] if not p.block_constant[int(b)]]
            cols = np.concatenate([np.arange(3 * pos[b], 3 * pos[b] + 3) for b in blocks])
            H[np.ix_(cols, cols)] += J.T @ J
    C = np.linalg.inv(H)
    fblocks = [b for b in active if p.block_group[b] != 0]
    pick = [fblocks[0], fblocks[7], fblocks[len(fblocks) // 2], fblocks[-1]]
    got = _cov_matrix(s, pick)
    idx = np.concatenate([np.arange(3 * pos[b], 3 * pos[b] + 3) for b in pick])
    want = C[np.ix_(idx, idx)]
    assert np.abs(got - want).max() <= 1e-7 * np.abs(want).max()
    lm = next(b for b in active if p.block_group[b] == 0)
    with pytest.raises(BsnlsError) as e:
        s.covariance([(lm, lm)])
    assert e.value.code == cabi.ERR_UNSUPPORTED
    # through the graph interface (UUID requests, GpuHashGraph::getCovariance)
    g.set_values(x)
    via_graph = g.covariance([(pick[0], pick[1])])[0]
    assert np.abs(via_graph - got[0:3, 3:6]).max() <= 1e-9 * np.abs(want).max()
