// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.h). Cost functors of the reference, restated.
#pragma once
#include <algorithm>
#include <limits>
#include "oracle_math.h"
#include "../include/bsnls.h"

namespace orc {

// ---------------------------------------------------------------------------------------------------
// fuse_constraints::NormalPriorOrientation3DCostFunctor::operator()
// (dependencies/fuse/fuse_constraints/include/fuse_constraints/normal_prior_orientation_3d_cost_functor.h:93-122)
// with A = identity (as used inside the pose / imu-state priors) — caller scales.
// ---------------------------------------------------------------------------------------------------
template <typename T> inline void PriorOrientationDelta(const T* q, const double* b, T* res3) {
  T observation_inverse[4] = {T(b[0]), T(-b[1]), T(-b[2]), T(-b[3])};
  T difference[4];
  QuaternionProduct(observation_inverse, q, difference);
  QuaternionToAngleAxis(difference, res3);
}
// fuse_constraints::NormalDeltaOrientation3DCostFunctor::operator()
// (.../normal_delta_orientation_3d_cost_functor.h:92-123), A applied by caller.
template <typename T> inline void DeltaOrientation(const T* q1, const T* q2, const double* b, T* res3) {
  T q1_inverse[4] = {q1[0], -q1[1], -q1[2], -q1[3]};
  T observation_inverse[4] = {T(b[0]), T(-b[1]), T(-b[2]), T(-b[3])};
  T difference[4];
  QuaternionProduct(q1_inverse, q2, difference);
  T error[4];
  QuaternionProduct(observation_inverse, difference, error);
  QuaternionToAngleAxis(error, res3);
}
// residual_map.applyOnTheLeft(A): r <- A r, A row-major n x n
template <typename T, int NR> inline void ApplyLeft(const double* A, T* r) {
  T tmp[NR];
  for (int i = 0; i < NR; ++i) {
    T acc = A[i * NR] * r[0];
    for (int j = 1; j < NR; ++j) acc = acc + A[i * NR + j] * r[j];
    tmp[i] = acc;
  }
  for (int i = 0; i < NR; ++i) r[i] = tmp[i];
}

// RelativeOrientation3DStampedConstraint: blocks q1,q2; consts delta[4], A[9]
struct RelOrientationFunctor {
  static constexpr int kRes = 3; static constexpr int kBlocks = 2;
  static const int* sizes() { static const int s[] = {4, 4}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    DeltaOrientation(p[0], p[1], c, r);
    ApplyLeft<T, 3>(c + 4, r);
  }
};

// fuse_constraints::NormalPriorOrientation3DCostFunctor::operator() (normal_prior_orientation_3d_cost_functor.h:92-125):
// AbsoluteOrientation3DStampedConstraint; blocks q; consts mean[4], A[9]
struct AbsOrientationFunctor {
  static constexpr int kRes = 3; static constexpr int kBlocks = 1;
  static const int* sizes() { static const int s[] = {4}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    PriorOrientationDelta(p[0], c, r);
    ApplyLeft<T, 3>(c + 4, r);
  }
};
// bs_constraints::GravityAlignmentCostFunctor::operator() (bs_constraints/global/gravity_alignment_cost_functor.h:53-66):
// blocks q_WORLD_BASELINK; consts gravity_in_baselink[3], A[4] (2x2 row-major)
struct GravityAlignmentFunctor {
  static constexpr int kRes = 2; static constexpr int kBlocks = 1;
  static const int* sizes() { static const int s[] = {4}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    const T g[3] = {T(c[0]), T(c[1]), T(c[2])}; T w[3];
    QuaternionRotatePoint(p[0], g, w);
    r[0] = c[3] * w[0] + c[4] * w[1]; r[1] = c[5] * w[0] + c[6] * w[1];
  }
};

// fuse_constraints::NormalDeltaPose3DCostFunctor::operator()
// (.../normal_delta_pose_3d_cost_functor.h:107-148): blocks p1,q1,p2,q2; consts b[7], A[36]
template <typename T>
inline void DeltaPose(const double* c, const T* p1, const T* q1, const T* p2, const T* q2, T* r) {
  T q1_inverse[4] = {q1[0], -q1[1], -q1[2], -q1[3]};
  T position_delta[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  T rotated[3];
  QuaternionRotatePoint(q1_inverse, position_delta, rotated);
  r[0] = rotated[0] - c[0]; r[1] = rotated[1] - c[1]; r[2] = rotated[2] - c[2];
  DeltaOrientation(q1, q2, c + 3, r + 3);
  ApplyLeft<T, 6>(c + 7, r);
}
struct RelPoseFunctor {
  static constexpr int kRes = 6; static constexpr int kBlocks = 4;
  static const int* sizes() { static const int s[] = {3, 4, 3, 4}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    DeltaPose(c, p[0], p[1], p[2], p[3], r);
  }
};
// bs_constraints::DeltaPose3DWithExtrinsicsCostFunctor
// (bs_constraints/include/bs_constraints/relative_pose/delta_pose_3d_with_extrinsics_cost_functor.h:65-109)
template <typename T>
inline void ToSensorFrame(const T* p_WB, const T* o_WB, const T* p_BS, const T* o_BS, T* p_WS, T* o_WS) {
  QuaternionProduct(o_WB, o_BS, o_WS);
  T t[3];
  QuaternionRotatePoint(o_WB, p_BS, t);
  p_WS[0] = t[0] + p_WB[0]; p_WS[1] = t[1] + p_WB[1]; p_WS[2] = t[2] + p_WB[2];
}
struct RelPoseExtFunctor {
  static constexpr int kRes = 6; static constexpr int kBlocks = 6;
  static const int* sizes() { static const int s[] = {3, 4, 3, 4, 3, 4}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    T p1[3], o1[4], p2[3], o2[4];
    ToSensorFrame(p[0], p[1], p[4], p[5], p1, o1);
    ToSensorFrame(p[2], p[3], p[4], p[5], p2, o2);
    DeltaPose(c, p1, o1, p2, o2, r);
  }
};
// fuse_constraints::NormalPriorPose3DCostFunctor (.../normal_prior_pose_3d_cost_functor.h:86-123)
struct AbsPoseFunctor {
  static constexpr int kRes = 6; static constexpr int kBlocks = 2;
  static const int* sizes() { static const int s[] = {3, 4}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    r[0] = p[0][0] - c[0]; r[1] = p[0][1] - c[1]; r[2] = p[0][2] - c[2];
    PriorOrientationDelta(p[1], c + 3, r + 3);
    ApplyLeft<T, 6>(c + 7, r);
  }
};
// bs_constraints::NormalPriorImuState3DCostFunctor
// (bs_constraints/include/bs_constraints/inertial/normal_prior_imu_state_3d_cost_functor.h:54-90)
struct ImuPriorFunctor {
  static constexpr int kRes = 15; static constexpr int kBlocks = 5;
  static const int* sizes() { static const int s[] = {4, 3, 3, 3, 3}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    PriorOrientationDelta(p[0], c, r);
    for (int b = 1; b < 5; ++b)
      for (int k = 0; k < 3; ++k) r[3 * b + k] = p[b][k] - c[4 + 3 * (b - 1) + k];
    ApplyLeft<T, 15>(c + 16, r);
  }
};
// bs_constraints::NormalDeltaImuState3DCostFunctor::operator()
// (bs_constraints/include/bs_constraints/inertial/normal_delta_imu_state_3d_cost_functor.h:59-141);
// DeltaQ: bs_common/include/bs_common/utils.h:28-38; GRAVITY_WORLD: utils.h:20-24.
struct ImuDeltaFunctor {
  static constexpr int kRes = 15; static constexpr int kBlocks = 10;
  static const int* sizes() { static const int s[] = {4, 3, 3, 3, 3, 4, 3, 3, 3, 3}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    const T* q_i = p[0]; const T* p_i = p[1]; const T* v_i = p[2]; const T* bg_i = p[3]; const T* ba_i = p[4];
    const T* q_j = p[5]; const T* p_j = p[6]; const T* v_j = p[7]; const T* bg_j = p[8]; const T* ba_j = p[9];
    const double dt = c[0];
    const double* dq = c + 1; const double* dp = c + 5; const double* dv = c + 8;
    const double* dq_dbg = c + 11; const double* dp_dbg = c + 20; const double* dp_dba = c + 29;
    const double* dv_dbg = c + 38; const double* dv_dba = c + 47;
    const double* bg0 = c + 56; const double* ba0 = c + 59; const double* A = c + 62;
    const double G[3] = {0.0, 0.0, -9.80665};
    T dbg[3], dba[3];
    for (int k = 0; k < 3; ++k) { dbg[k] = bg_i[k] - bg0[k]; dba[k] = ba_i[k] - ba0[k]; }
    T q_tmp[3]; Mat3Vec(dq_dbg, dbg, q_tmp);
    T dQ[4] = {T(1.0), q_tmp[0] / 2.0, q_tmp[1] / 2.0, q_tmp[2] / 2.0};
    T dqT[4] = {T(dq[0]), T(dq[1]), T(dq[2]), T(dq[3])};
    T q_corrected[4]; EigenQuatProduct(dqT, dQ, q_corrected);
    T t1[3], t2[3], p_corrected[3], v_corrected[3];
    Mat3Vec(dp_dbg, dbg, t1); Mat3Vec(dp_dba, dba, t2);
    for (int k = 0; k < 3; ++k) p_corrected[k] = (dp[k] + t1[k]) + t2[k];
    Mat3Vec(dv_dbg, dbg, t1); Mat3Vec(dv_dba, dba, t2);
    for (int k = 0; k < 3; ++k) v_corrected[k] = (dv[k] + t1[k]) + t2[k];
    // res_q = 2 * (q_corrected.inverse() * (q_i.inverse() * q_j)).vec()
    T qi_inv[4], qc_inv[4], qij[4], qe[4];
    EigenQuatInverse(q_i, qi_inv); EigenQuatProduct(qi_inv, q_j, qij);
    EigenQuatInverse(q_corrected, qc_inv); EigenQuatProduct(qc_inv, qij, qe);
    r[0] = 2.0 * qe[1]; r[1] = 2.0 * qe[2]; r[2] = 2.0 * qe[3];
    // res_p = q_i.conjugate() * (p_j - p_i - dt*v_i - 0.5*dt*dt*G) - p_corrected
    T qi_conj[4] = {q_i[0], -q_i[1], -q_i[2], -q_i[3]};
    T arg[3], rot[3];
    for (int k = 0; k < 3; ++k) arg[k] = ((p_j[k] - p_i[k]) - dt * v_i[k]) - ((0.5 * dt) * dt) * G[k];
    EigenQuatRotate(qi_conj, arg, rot);
    for (int k = 0; k < 3; ++k) r[3 + k] = rot[k] - p_corrected[k];
    for (int k = 0; k < 3; ++k) arg[k] = (v_j[k] - v_i[k]) - dt * G[k];
    EigenQuatRotate(qi_conj, arg, rot);
    for (int k = 0; k < 3; ++k) r[6 + k] = rot[k] - v_corrected[k];
    for (int k = 0; k < 3; ++k) { r[9 + k] = bg_j[k] - bg_i[k]; r[12 + k] = ba_j[k] - ba_i[k]; }
    ApplyLeft<T, 15>(A, r);
  }
};
// bs_constraints::AbsoluteConstraint<T> / RelativeConstraint<T> on 3-vectors
// (bs_constraints/include/bs_constraints/global/absolute_constraint.h:14-28 -> fuse NormalPrior / NormalDelta)
struct AbsVec3Functor {
  static constexpr int kRes = 3; static constexpr int kBlocks = 1;
  static const int* sizes() { static const int s[] = {3}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    for (int k = 0; k < 3; ++k) r[k] = p[0][k] - c[k];
    ApplyLeft<T, 3>(c + 3, r);
  }
};
struct RelVec3Functor {
  static constexpr int kRes = 3; static constexpr int kBlocks = 2;
  static const int* sizes() { static const int s[] = {3, 3}; return s; }
  template <typename T> static void eval(const double* c, const T* const* p, T* r) {
    for (int k = 0; k < 3; ++k) r[k] = (p[1][k] - p[0][k]) - c[k];
    ApplyLeft<T, 3>(c + 3, r);
  }
};

// ceres::AutoDiffCostFunction<Functor, R, N0, N1, ...>::Evaluate: one forward pass with Jet<sum Ni>.
// jac[b] (may be null) is row-major kRes x size_b.
template <class F, int NT>
inline void AutoDiffEvaluate(const double* consts, const double* const* params, double* residual,
                             double** jac) {
  const int* sz = F::sizes();
  if (!jac) {
    F::template eval<double>(consts, params, residual);
    return;
  }
  using J = Jet<NT>;
  J store[NT];
  const J* pp[F::kBlocks];
  int off = 0;
  for (int b = 0; b < F::kBlocks; ++b) {
    pp[b] = store + off;
    for (int k = 0; k < sz[b]; ++k) store[off + k] = J(params[b][k], off + k);
    off += sz[b];
  }
  J res[F::kRes];
  F::template eval<J>(consts, pp, res);
  off = 0;
  for (int b = 0; b < F::kBlocks; ++b) {
    for (int i = 0; i < F::kRes; ++i) {
      if (b == 0) residual[i] = res[i].a;
      if (jac[b]) for (int k = 0; k < sz[b]; ++k) jac[b][i * sz[b] + k] = res[i].v[off + k];
    }
    off += sz[b];
  }
}

// ---------------------------------------------------------------------------------------------------
// bs_constraints::EuclideanReprojection::Evaluate
// (bs_constraints/include/bs_constraints/visual/euclidean_reprojection_function.h:66-172) with the helpers
// DImageProjectionDPoint / DPointRotationDPoint (bs_constraints/src/jacobians.cpp:181-214) and the
// file-scope `point_transformation` lambda (:15-24) whose forward difference (EPSILON = 1e-8) is the
// quaternion Jacobian. Eigen evaluation order is kept: 3-term dot products are a0 + (a1 + a2); the build is
// assumed to be the default x86-64 one (no FMA contraction) — this file is compiled with -ffp-contract=off.
// cam: fx, fy, cx, cy, T_cam_baselink (3x4 row-major), w.
// ---------------------------------------------------------------------------------------------------
inline double dot3(double a0, double b0, double a1, double b1, double a2, double b2) {
  return a0 * b0 + (a1 * b1 + a2 * b2);
}
inline void QuatToRotEigen(double w, double x, double y, double z, double* R /*row-major*/) {
  const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1.0 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1.0 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1.0 - (txx + tyy);
}
// point_transformation(q_coeffs, t, P): R = q.normalized().toRotationMatrix(); (R^T P) - (R^T t)
inline void PointTransformation(const double* q, const double* t, const double* P, double* out) {
  // Eigen::Quaternion::normalized() -> coeffs (x,y,z,w).normalized(): z = squaredNorm; coeffs / sqrt(z).
  // SSE2 packet reduction of (x2,y2,z2,w2): (x2+z2) + (y2+w2).
  const double n2 = (q[1] * q[1] + q[3] * q[3]) + (q[2] * q[2] + q[0] * q[0]);
  double w = q[0], x = q[1], y = q[2], z = q[3];
  if (n2 > 0.0) { const double n = std::sqrt(n2); w = w / n; x = x / n; y = y / n; z = z / n; }
  double R[9]; QuatToRotEigen(w, x, y, z, R);
  for (int i = 0; i < 3; ++i) {
    const double a = dot3(R[i], P[0], R[3 + i], P[1], R[6 + i], P[2]);
    const double b = dot3(R[i], t[0], R[3 + i], t[1], R[6 + i], t[2]);
    out[i] = a - b;
  }
}
inline void ReprojectionEvaluate(const double* cam, const double* pixel, const double* q, const double* t,
                                 const double* P, double* residual, double* Jq /*2x4*/, double* Jt /*2x3*/,
                                 double* JP /*2x3*/) {
  const double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
  const double* T = cam + 4;  // 3x4 row-major
  const double w = cam[16];
  double R[9]; QuatToRotEigen(q[0], q[1], q[2], q[3], R);  // R_WORLD_BASELINK (not normalised)
  // P_BASELINK = (R^T P) - (R^T t)
  double Pb[3];
  for (int i = 0; i < 3; ++i)
    Pb[i] = dot3(R[i], P[0], R[3 + i], P[1], R[6 + i], P[2]) - dot3(R[i], t[0], R[3 + i], t[1], R[6 + i], t[2]);
  // P_CAMERA = R_cb * P_b + t_cb
  double Pc[3];
  for (int i = 0; i < 3; ++i) Pc[i] = dot3(T[4 * i], Pb[0], T[4 * i + 1], Pb[1], T[4 * i + 2], Pb[2]) + T[4 * i + 3];
  // (K * P_CAMERA).hnormalized()
  const double k0 = dot3(fx, Pc[0], 0.0, Pc[1], cx, Pc[2]);
  const double k1 = dot3(0.0, Pc[0], fy, Pc[1], cy, Pc[2]);
  const double k2 = dot3(0.0, Pc[0], 0.0, Pc[1], 1.0, Pc[2]);
  const double u = k0 / k2, v = k1 / k2;
  // E = W * (pixel - reprojection), W = w*I2
  const double e0 = pixel[0] - u, e1 = pixel[1] - v;
  residual[0] = w * e0 + 0.0 * e1;
  residual[1] = 0.0 * e0 + w * e1;
  if (!Jq && !Jt && !JP) return;
  // d_E_d_P_CAMERA (2x3)
  const double z = Pc[2], z2 = z * z;
  const double dE[6] = {fx / z, 0.0, -fx * Pc[0] / z2, 0.0, fy / z, -fy * Pc[1] / z2};
  // A = d_E_d_P_CAMERA * R_cb (2x3)
  double A[6];
  for (int i = 0; i < 2; ++i)
    for (int j = 0; j < 3; ++j) A[3 * i + j] = dot3(dE[3 * i], T[j], dE[3 * i + 1], T[4 + j], dE[3 * i + 2], T[8 + j]);
  auto left_minus_W = [&](double* J, int cols) {  // J <- (-W) * J
    for (int j = 0; j < cols; ++j) {
      const double a = J[j], b = J[cols + j];
      J[j] = (-w) * a + (-0.0) * b;
      J[cols + j] = (-0.0) * a + (-w) * b;
    }
  };
  if (Jq) {
    double res[3]; PointTransformation(q, t, P, res);
    double D[12];  // 3x4, column i = finite difference
    for (int i = 0; i < 4; ++i) {
      double qp[4] = {q[0] + 0.0, q[1] + 0.0, q[2] + 0.0, q[3] + 0.0};
      qp[i] = q[i] + 1e-8;
      double rp[3]; PointTransformation(qp, t, P, rp);
      for (int k = 0; k < 3; ++k) D[4 * k + i] = (rp[k] - res[k]) / 1e-8;
    }
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 4; ++j) Jq[4 * i + j] = dot3(A[3 * i], D[j], A[3 * i + 1], D[4 + j], A[3 * i + 2], D[8 + j]);
    left_minus_W(Jq, 4);
  }
  if (Jt) {  // d_P_BASELINK_d_t = -R^T
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 3; ++j) Jt[3 * i + j] = dot3(A[3 * i], -R[3 * j], A[3 * i + 1], -R[3 * j + 1], A[3 * i + 2], -R[3 * j + 2]);
    left_minus_W(Jt, 3);
  }
  if (JP) {  // d_P_BASELINK_d_P = R^T
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 3; ++j) JP[3 * i + j] = dot3(A[3 * i], R[3 * j], A[3 * i + 1], R[3 * j + 1], A[3 * i + 2], R[3 * j + 2]);
    left_minus_W(JP, 3);
  }
}

// ---------------------------------------------------------------------------------------------------
// fuse_variables::Orientation3DLocalParameterization (orientation_3d_stamped.h:91-142)
// ---------------------------------------------------------------------------------------------------
inline void QuatPlus(const double* x, const double* delta, double* out) {
  double qd[4]; AngleAxisToQuaternion(delta, qd); QuaternionProduct(x, qd, out);
}
inline void QuatPlusJacobian(const double* x, double* J /*4x3 row-major*/) {
  const double x0 = x[0] / 2, x1 = x[1] / 2, x2 = x[2] / 2, x3 = x[3] / 2;
  J[0] = -x1; J[1] = -x2; J[2] = -x3;
  J[3] = x0;  J[4] = -x3; J[5] = x2;
  J[6] = x3;  J[7] = x0;  J[8] = -x1;
  J[9] = -x2; J[10] = x1; J[11] = x0;
}
inline void QuatMinus(const double* x1, const double* x2, double* delta) {
  double inv[4] = {x1[0], -x1[1], -x1[2], -x1[3]}, qd[4];
  QuaternionProduct(inv, x2, qd); QuaternionToAngleAxis(qd, delta);
}

// fuse_constraints::MarginalCostFunction::Evaluate (fuse_constraints/src/marginal_cost_function.cpp:58-113):
// residual = b + sum_i A_i Minus(x_bar_i, x_i); jacobian_i = A_i * MinusJacobian(x_i) for a quaternion block
// (orientation_3d_stamped.h:130-142), A_i for a vector block. consts as documented at BSNLS_FACTOR_MARGINAL.
inline void MarginalEvaluate(const double* c, const double* const* params, double* residual, double** jac) {
  const int nb = (int)c[0];
  const double* bvec = c + 11; const double* xbar = c + 26; const double* A = c + 58;
  for (int i = 0; i < BSNLS_MARGINAL_MAX_ROWS; ++i) residual[i] = bvec[i];
  int off = 0;
  for (int b = 0; b < nb; ++b) {
    const int sz = (int)c[1 + b];
    double d[3];
    if (sz == 4) QuatMinus(xbar + off, params[b], d);
    else for (int k = 0; k < 3; ++k) d[k] = params[b][k] - xbar[off + k];
    for (int i = 0; i < BSNLS_MARGINAL_MAX_ROWS; ++i) {
      const double* a = A + i * 30 + 3 * b;
      residual[i] += a[0] * d[0] + a[1] * d[1] + a[2] * d[2];
    }
    if (jac && jac[b]) {
      if (sz == 4) {
        const double* x = params[b];
        const double x0 = 2 * x[0], x1 = 2 * x[1], x2 = 2 * x[2], x3 = 2 * x[3];
        const double Jm[12] = {-x1, x0, x3, -x2,  -x2, -x3, x0, x1,  -x3, x2, -x1, x0};   // 3x4 row-major
        for (int i = 0; i < BSNLS_MARGINAL_MAX_ROWS; ++i) {
          const double* a = A + i * 30 + 3 * b;
          for (int k = 0; k < 4; ++k) jac[b][i * 4 + k] = a[0] * Jm[k] + a[1] * Jm[4 + k] + a[2] * Jm[8 + k];
        }
      } else {
        for (int i = 0; i < BSNLS_MARGINAL_MAX_ROWS; ++i) for (int k = 0; k < 3; ++k) jac[b][i * 3 + k] = A[i * 30 + 3 * b + k];
      }
    }
    off += sz;
  }
}

// ceres::CauchyLoss / HuberLoss ::Evaluate (ceres/loss_function.cc); fuse wrappers:
// dependencies/fuse/fuse_loss/src/cauchy_loss.cpp:65-68, huber_loss.cpp.
inline void LossEvaluate(int type, double a, double s, double rho[3]) {
  if (type == BSNLS_LOSS_CAUCHY) {
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, inv = 1.0 / sum;
    rho[0] = b * std::log(sum); rho[1] = std::max(std::numeric_limits<double>::min(), inv);
    rho[2] = -c * (inv * inv);
  } else if (type == BSNLS_LOSS_HUBER) {
    const double b = a * a;
    if (s > b) { const double r = std::sqrt(s); rho[0] = 2.0 * a * r - b;
      rho[1] = std::max(std::numeric_limits<double>::min(), a / r); rho[2] = -rho[1] / (2.0 * s); }
    else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
  } else if (type == BSNLS_LOSS_SOFTLONE) {   // ceres::SoftLOneLoss
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, tmp = std::sqrt(sum);
    rho[0] = 2.0 * b * (tmp - 1.0); rho[1] = std::max(std::numeric_limits<double>::min(), 1.0 / tmp); rho[2] = -(c * rho[1]) / (2.0 * sum);
  } else if (type == BSNLS_LOSS_ARCTAN) {     // ceres::ArctanLoss
    const double b = 1.0 / (a * a);
    const double sum = 1.0 + s * s * b, inv = 1.0 / sum;
    rho[0] = a * std::atan2(s, a); rho[1] = std::max(std::numeric_limits<double>::min(), inv); rho[2] = -2.0 * s * b * (inv * inv);
  } else if (type == BSNLS_LOSS_TUKEY) {      // ceres::TukeyLoss (Ceres 2.x scaling)
    const double b = a * a;
    if (s <= b) { const double v = 1.0 - s / b, v2 = v * v; rho[0] = b / 3.0 * (1.0 - v2 * v); rho[1] = v2; rho[2] = -2.0 / b * v; }
    else { rho[0] = b / 3.0; rho[1] = 0.0; rho[2] = 0.0; }
  } else if (type == BSNLS_LOSS_DCS) {        // fuse_loss/src/loss_function.cpp:44-62
    if (s > a) { const double inv = 1.0 / (a + s), scale = 2.0 * a * inv; rho[0] = a * (3.0 * s - a) * inv; rho[1] = scale * scale; rho[2] = -2.0 * inv * rho[1]; }
    else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
  } else if (type == BSNLS_LOSS_FAIR) {       // fuse_loss/src/loss_function.cpp:64-73 (b = a^2)
    const double r = std::sqrt(s), ra = r / a, sum = 1.0 + ra;
    rho[0] = 2.0 * (a * a) * (ra - std::log(sum)); rho[1] = 1.0 / sum; rho[2] = r == 0.0 ? -std::numeric_limits<double>::max() : -0.5 / (a * r * sum * sum);
  } else if (type == BSNLS_LOSS_GEMAN_MCCLURE) {   // fuse_loss/src/loss_function.cpp:75-84
    const double b = a * a, sum = b + s, inv = 1.0 / sum, scale = b * inv;
    rho[0] = s * scale; rho[1] = scale * scale; rho[2] = -2.0 * inv * rho[1];
  } else if (type == BSNLS_LOSS_WELSCH) {     // fuse_loss/src/loss_function.cpp:86-93
    const double b = a * a, c = -1.0 / b, e = std::exp(s * c);
    rho[0] = b * (1.0 - e); rho[1] = e; rho[2] = c * e;
  } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}

}  // namespace orc
