// Device-side factor math for the fuse-graph LM solve (sm_100a, FP64).
//
// Reprojection: bs_constraints::EuclideanReprojection::Evaluate
//   (bs_constraints/include/bs_constraints/visual/euclidean_reprojection_function.h:66-172) including its
//   forward-difference quaternion Jacobian (EPSILON = 1e-8, :13-24,124-143). The finite difference divides
//   rounding noise by 1e-8, so that part is written with explicit non-fused IEEE operations (__dmul_rn /
//   __dadd_rn / __dsub_rn, IEEE sqrt and division) in Eigen's evaluation order: the reference's default x86-64
//   build has no FMA contraction, and the result is then bit-reproducible.
// Autodiff factors (IMU, relative pose, priors): the reference differentiates them with ceres::Jet
//   (AutoDiffCostFunction). Here ONE WARP evaluates one factor: lane k carries the dual part for ambient
//   parameter k (<= 32 ambient parameters for every factor type), i.e. forward mode with a 1-wide dual number
//   per lane, which yields the same derivatives as the 32-wide Jet.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace bsnls {

// ---------------------------------------------------------------------------------------------------
// Exact (unfused) helpers for the finite-difference path
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ double xmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double xadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double xsub(double a, double b) { return __dsub_rn(a, b); }
// Eigen fixed-size 3-term dot product: a0*b0 + (a1*b1 + a2*b2)  (Eigen/src/Core/Redux.h redux_novec_unroller)
__device__ __forceinline__ double xdot3(double a0, double b0, double a1, double b1, double a2, double b2) {
  return xadd(xmul(a0, b0), xadd(xmul(a1, b1), xmul(a2, b2)));
}
// Eigen::Quaternion::toRotationMatrix (Eigen/src/Geometry/Quaternion.h), row-major
__device__ __forceinline__ void quat_to_rot_exact(double w, double x, double y, double z, double* R) {
  const double tx = xmul(2.0, x), ty = xmul(2.0, y), tz = xmul(2.0, z);
  const double twx = xmul(tx, w), twy = xmul(ty, w), twz = xmul(tz, w);
  const double txx = xmul(tx, x), txy = xmul(ty, x), txz = xmul(tz, x);
  const double tyy = xmul(ty, y), tyz = xmul(tz, y), tzz = xmul(tz, z);
  R[0] = xsub(1.0, xadd(tyy, tzz)); R[1] = xsub(txy, twz); R[2] = xadd(txz, twy);
  R[3] = xadd(txy, twz); R[4] = xsub(1.0, xadd(txx, tzz)); R[5] = xsub(tyz, twx);
  R[6] = xsub(txz, twy); R[7] = xadd(tyz, twx); R[8] = xsub(1.0, xadd(txx, tyy));
}
// the file-scope `point_transformation` lambda (euclidean_reprojection_function.h:15-24):
// R = q.normalized().toRotationMatrix(); (R^T P) - (R^T t)
__device__ __forceinline__ void point_transformation_exact(const double* q, const double* t, const double* P,
                                                           double* out) {
  // coeffs().squaredNorm() over (x,y,z,w), SSE2 packet order: (x2 + z2) + (y2 + w2)
  const double n2 = xadd(xadd(xmul(q[1], q[1]), xmul(q[3], q[3])), xadd(xmul(q[2], q[2]), xmul(q[0], q[0])));
  double w = q[0], x = q[1], y = q[2], z = q[3];
  if (n2 > 0.0) { const double n = __dsqrt_rn(n2); w = __ddiv_rn(w, n); x = __ddiv_rn(x, n); y = __ddiv_rn(y, n); z = __ddiv_rn(z, n); }
  double R[9]; quat_to_rot_exact(w, x, y, z, R);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double a = xdot3(R[i], P[0], R[3 + i], P[1], R[6 + i], P[2]);
    const double b = xdot3(R[i], t[0], R[3 + i], t[1], R[6 + i], t[2]);
    out[i] = xsub(a, b);
  }
}

struct CameraModel {  // one row of bsnls_problem::cameras
  double fx, fy, cx, cy; double T[12]; double w; double loss_type; double loss_a; double pad;
};

// ceres / fuse loss functions: rho[0..2] = rho(s), rho'(s), rho''(s). type = BSNLS_LOSS_* (include/bsnls.h)
__device__ __forceinline__ void loss_evaluate(int type, double a, double s, double* rho) {
  if (type == 1) {  // CauchyLoss
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, inv = 1.0 / sum;
    rho[0] = b * log(sum); rho[1] = fmax(2.2250738585072014e-308, inv); rho[2] = -c * (inv * inv);
  } else if (type == 2) {  // HuberLoss
    const double b = a * a;
    if (s > b) { const double r = sqrt(s); rho[0] = 2.0 * a * r - b; rho[1] = fmax(2.2250738585072014e-308, a / r); rho[2] = -rho[1] / (2.0 * s); }
    else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
  } else if (type == BSNLS_LOSS_SOFTLONE) {   // ceres::SoftLOneLoss
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, tmp = sqrt(sum);
    rho[0] = 2.0 * b * (tmp - 1.0); rho[1] = fmax(2.2250738585072014e-308, 1.0 / tmp); rho[2] = -(c * rho[1]) / (2.0 * sum);
  } else if (type == BSNLS_LOSS_ARCTAN) {     // ceres::ArctanLoss
    const double b = 1.0 / (a * a);
    const double sum = 1.0 + s * s * b, inv = 1.0 / sum;
    rho[0] = a * atan2(s, a); rho[1] = fmax(2.2250738585072014e-308, inv); rho[2] = -2.0 * s * b * (inv * inv);
  } else if (type == BSNLS_LOSS_TUKEY) {      // ceres::TukeyLoss (Ceres 2.x scaling)
    const double b = a * a;
    if (s <= b) { const double v = 1.0 - s / b, v2 = v * v; rho[0] = b / 3.0 * (1.0 - v2 * v); rho[1] = v2; rho[2] = -2.0 / b * v; }
    else { rho[0] = b / 3.0; rho[1] = 0.0; rho[2] = 0.0; }
  } else if (type == BSNLS_LOSS_DCS) {        // fuse_loss/src/loss_function.cpp:44-62
    if (s > a) { const double inv = 1.0 / (a + s), scale = 2.0 * a * inv; rho[0] = a * (3.0 * s - a) * inv; rho[1] = scale * scale; rho[2] = -2.0 * inv * rho[1]; }
    else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
  } else if (type == BSNLS_LOSS_FAIR) {       // fuse_loss/src/loss_function.cpp:64-73 (b = a^2)
    const double r = sqrt(s), ra = r / a, sum = 1.0 + ra;
    rho[0] = 2.0 * (a * a) * (ra - log(sum)); rho[1] = 1.0 / sum; rho[2] = r == 0.0 ? -1.7976931348623157e308 : -0.5 / (a * r * sum * sum);
  } else if (type == BSNLS_LOSS_GEMAN_MCCLURE) {   // fuse_loss/src/loss_function.cpp:75-84
    const double b = a * a, sum = b + s, inv = 1.0 / sum, scale = b * inv;
    rho[0] = s * scale; rho[1] = scale * scale; rho[2] = -2.0 * inv * rho[1];
  } else if (type == BSNLS_LOSS_WELSCH) {     // fuse_loss/src/loss_function.cpp:86-93
    const double b = a * a, c = -1.0 / b, e = exp(s * c);
    rho[0] = b * (1.0 - e); rho[1] = e; rho[2] = c * e;
  } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}
// ceres Corrector (corrector.cc; restated in-tree at fuse_constraints/src/marginalize_variables.cpp:361-396):
// returns sqrt(rho'), residual scaling and alpha/s for the Jacobian correction.
__device__ __forceinline__ void corrector(double s, const double* rho, double* sqrt_rho1, double* res_scale,
                                          double* alpha_sq_norm) {
  *sqrt_rho1 = sqrt(rho[1]);
  if (s == 0.0 || rho[2] <= 0.0) { *res_scale = *sqrt_rho1; *alpha_sq_norm = 0.0; return; }
  const double D = 1.0 + 2.0 * s * rho[2] / rho[1];
  const double alpha = 1.0 - sqrt(D);
  *res_scale = *sqrt_rho1 / (1.0 - alpha); *alpha_sq_norm = alpha / s;
}

// Orientation3DLocalParameterization::ComputeJacobian (orientation_3d_stamped.h:106-115), 4x3 row-major
__device__ __forceinline__ void quat_plus_jacobian(const double* x, double* J) {
  const double x0 = x[0] / 2, x1 = x[1] / 2, x2 = x[2] / 2, x3 = x[3] / 2;
  J[0] = -x1; J[1] = -x2; J[2] = -x3; J[3] = x0; J[4] = -x3; J[5] = x2;
  J[6] = x3; J[7] = x0; J[8] = -x1; J[9] = -x2; J[10] = x1; J[11] = x0;
}

// Residual only (candidate-cost evaluation): r = W (pixel - pi(K (R_cb R^T (P - t) + t_cb)))
__device__ __forceinline__ void reprojection_residual(const CameraModel& cam, const double* pixel, const double* q,
                                                      const double* t, const double* P, double* r, double* R,
                                                      double* Pc) {
  quat_to_rot_exact(q[0], q[1], q[2], q[3], R);  // R_WORLD_BASELINK, not normalised (:74-75)
  double Pb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    Pb[i] = xsub(xdot3(R[i], P[0], R[3 + i], P[1], R[6 + i], P[2]), xdot3(R[i], t[0], R[3 + i], t[1], R[6 + i], t[2]));
#pragma unroll
  for (int i = 0; i < 3; ++i)
    Pc[i] = xadd(xdot3(cam.T[4 * i], Pb[0], cam.T[4 * i + 1], Pb[1], cam.T[4 * i + 2], Pb[2]), cam.T[4 * i + 3]);
  // K = [fx 0 cx; 0 fy cy; 0 0 1] and W = w I: the reference's products with the structural zeros add +-0 and are
  // dropped here (identical values; only the sign of an exact zero could differ)
  const double k0 = xadd(xmul(cam.fx, Pc[0]), xmul(cam.cx, Pc[2]));
  const double k1 = xadd(xmul(cam.fy, Pc[1]), xmul(cam.cy, Pc[2]));
  const double u = __ddiv_rn(k0, Pc[2]), v = __ddiv_rn(k1, Pc[2]);
  r[0] = xmul(cam.w, xsub(pixel[0], u));
  r[1] = xmul(cam.w, xsub(pixel[1], v));
}

// ---------------------------------------------------------------------------------------------------
// Pose table. Everything in the reprojection Jacobian that depends on the pose alone is computed once per pose and
// pass instead of once per observation: the rotation of the raw quaternion (residual, t and P blocks) and the five
// rotations of point_transformation (normalised q and q + 1e-8 e_i, the forward difference of
// euclidean_reprojection_function.h:15-24, 77-90), each with R^T t. Same operations in the same order as
// reprojection_evaluate, so the results are bit-identical; only the redundancy across observations is gone.
// Layout: 6 sets x (R[9] row-major, R^T t [3]).
// ---------------------------------------------------------------------------------------------------
constexpr int kPoseTab = 72;
__device__ __forceinline__ void pose_table_set(const double* R, const double* t, double* out) {
#pragma unroll
  for (int k = 0; k < 9; ++k) out[k] = R[k];
#pragma unroll
  for (int i = 0; i < 3; ++i) out[9 + i] = xdot3(R[i], t[0], R[3 + i], t[1], R[6 + i], t[2]);
}
__device__ __forceinline__ void pose_table_build(const double* q, const double* t, double* tab) {
  double R[9];
  quat_to_rot_exact(q[0], q[1], q[2], q[3], R);
  pose_table_set(R, t, tab);
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    double qp[4] = {q[0], q[1], q[2], q[3]};
    if (k > 0) qp[k - 1] = xadd(q[k - 1], 1e-8);
    const double n2 = xadd(xadd(xmul(qp[1], qp[1]), xmul(qp[3], qp[3])), xadd(xmul(qp[2], qp[2]), xmul(qp[0], qp[0])));
    double w = qp[0], x = qp[1], y = qp[2], z = qp[3];
    if (n2 > 0.0) { const double n = __dsqrt_rn(n2); w = __ddiv_rn(w, n); x = __ddiv_rn(x, n); y = __ddiv_rn(y, n); z = __ddiv_rn(z, n); }
    quat_to_rot_exact(w, x, y, z, R);
    pose_table_set(R, t, tab + 12 * (k + 1));
  }
}
// x / 1e-8 correctly rounded (== __ddiv_rn(x, 1e-8)) without the general division sequence: RN(1 / 1e-8) = 1e8 exactly,
// two Markstein corrections with exact FMA residuals. |x| here is a difference of O(1..100) coordinates: no underflow.
__device__ __forceinline__ double fd_quotient(double x) {
  const double c = 1e-8, rc = 1e8;
  const double q0 = __dmul_rn(x, rc), r0 = __fma_rn(-c, q0, x), q1 = __fma_rn(r0, rc, q0), r1 = __fma_rn(-c, q1, x);
  return __fma_rn(r1, rc, q1);
}
// Full evaluation: residual + tangent Jacobian (2x3 wrt the q tangent, 2x3 wrt P), loss-corrected, with the pose-only
// parts read from the pose table (T = the pose's 72 doubles). Returns the cost contribution 0.5*rho(s). Jq/JP row-major
// 2x3. The block wrt t is not computed: the reference forms it as A * (-R) against A * R for P, and negation is exact
// through every step (products, the -W scaling, the loss correction), so Jt == -JP bit for bit.
__device__ __forceinline__ double reprojection_evaluate_tab(const CameraModel& cam, const double* pixel, const double* q,
                                                            const double* __restrict__ T, const double* P,
                                                            double* r, double* Jq, double* JP) {
  double R[9], Pc[3], Pb[3];
#pragma unroll
  for (int k = 0; k < 9; ++k) R[k] = T[k];
#pragma unroll
  for (int i = 0; i < 3; ++i) Pb[i] = xsub(xdot3(R[i], P[0], R[3 + i], P[1], R[6 + i], P[2]), T[9 + i]);
#pragma unroll
  for (int i = 0; i < 3; ++i)
    Pc[i] = xadd(xdot3(cam.T[4 * i], Pb[0], cam.T[4 * i + 1], Pb[1], cam.T[4 * i + 2], Pb[2]), cam.T[4 * i + 3]);
  {
    const double k0 = xadd(xmul(cam.fx, Pc[0]), xmul(cam.cx, Pc[2]));
    const double k1 = xadd(xmul(cam.fy, Pc[1]), xmul(cam.cy, Pc[2]));
    const double u = __ddiv_rn(k0, Pc[2]), v = __ddiv_rn(k1, Pc[2]);
    r[0] = xmul(cam.w, xsub(pixel[0], u));
    r[1] = xmul(cam.w, xsub(pixel[1], v));
  }
  const double z = Pc[2];
  // d_E_d_P_CAMERA = [fx/z 0 -fx X/z^2; 0 fy/z -fy Y/z^2]; A = d_E_d_P_CAMERA * R_cb without the structural-zero products
  // one reciprocal instead of four divisions: this block is analytic (not differenced), so a last-bit difference to the
  // reference's fx / z, fx X / z^2 is harmless (unlike the forward-difference block below, which is kept bit-exact)
  const double iz = 1.0 / z, iz2 = iz * iz;
  const double d00 = cam.fx * iz, d02 = -cam.fx * Pc[0] * iz2, d11 = cam.fy * iz, d12 = -cam.fy * Pc[1] * iz2;
  double A[6];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    A[j] = xadd(xmul(d00, cam.T[j]), xmul(d02, cam.T[8 + j]));
    A[3 + j] = xadd(xmul(d11, cam.T[4 + j]), xmul(d12, cam.T[8 + j]));
  }
  double res0[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) res0[k] = xsub(xdot3(T[12 + k], P[0], T[15 + k], P[1], T[18 + k], P[2]), T[21 + k]);
  double D[12];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double* Tp = T + 24 + 12 * i;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double rp = xsub(xdot3(Tp[k], P[0], Tp[3 + k], P[1], Tp[6 + k], P[2]), Tp[9 + k]);
      D[4 * k + i] = fd_quotient(xsub(rp, res0[k]));
    }
  }
  double J4[8];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      J4[4 * i + j] = xdot3(A[3 * i], D[j], A[3 * i + 1], D[4 + j], A[3 * i + 2], D[8 + j]);
  const double mw = -cam.w;
#pragma unroll
  for (int j = 0; j < 8; ++j) J4[j] = xmul(mw, J4[j]);   // applyOnTheLeft(-W), W = w I2
  double Pj[12]; quat_plus_jacobian(q, Pj);
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < 4; ++a) acc += J4[4 * i + a] * Pj[3 * a + k];
      Jq[3 * i + k] = acc;
    }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      JP[3 * i + j] = xmul(mw, xdot3(A[3 * i], R[3 * j], A[3 * i + 1], R[3 * j + 1], A[3 * i + 2], R[3 * j + 2]));
  const double s = r[0] * r[0] + r[1] * r[1];
  const int lt = static_cast<int>(cam.loss_type);
  if (lt == 0) return 0.5 * s;
  double rho[3]; loss_evaluate(lt, cam.loss_a, s, rho);
  double sr, rs, asn; corrector(s, rho, &sr, &rs, &asn);
  if (asn == 0.0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) { Jq[k] *= sr; JP[k] *= sr; }
  } else {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      double rtj = Jq[c] * r[0] + Jq[3 + c] * r[1];
      Jq[c] = sr * (Jq[c] - asn * r[0] * rtj); Jq[3 + c] = sr * (Jq[3 + c] - asn * r[1] * rtj);
      rtj = JP[c] * r[0] + JP[3 + c] * r[1];
      JP[c] = sr * (JP[c] - asn * r[0] * rtj); JP[3 + c] = sr * (JP[3 + c] - asn * r[1] * rtj);
    }
  }
  r[0] *= rs; r[1] *= rs;
  return 0.5 * rho[0];
}
__device__ __forceinline__ double reprojection_cost_only(const CameraModel& cam, const double* pixel,
                                                         const double* q, const double* t, const double* P) {
  double R[9], Pc[3], r[2];
  reprojection_residual(cam, pixel, q, t, P, r, R, Pc);
  const double s = r[0] * r[0] + r[1] * r[1];
  const int lt = static_cast<int>(cam.loss_type);
  if (lt == 0) return 0.5 * s;
  double rho[3]; loss_evaluate(lt, cam.loss_a, s, rho);
  return 0.5 * rho[0];
}

// ---------------------------------------------------------------------------------------------------
// 1-wide dual numbers (value + derivative along this lane's ambient parameter); ceres/jet.h rules.
// ---------------------------------------------------------------------------------------------------
struct Dual {
  double a, v;
  __device__ __forceinline__ Dual() : a(0.0), v(0.0) {}
  __device__ __forceinline__ Dual(double s) : a(s), v(0.0) {}  // NOLINT
  __device__ __forceinline__ Dual(double s, double d) : a(s), v(d) {}
};
__device__ __forceinline__ Dual operator+(const Dual& f, const Dual& g) { return Dual(f.a + g.a, f.v + g.v); }
__device__ __forceinline__ Dual operator-(const Dual& f, const Dual& g) { return Dual(f.a - g.a, f.v - g.v); }
__device__ __forceinline__ Dual operator-(const Dual& f) { return Dual(-f.a, -f.v); }
__device__ __forceinline__ Dual operator*(const Dual& f, const Dual& g) { return Dual(f.a * g.a, f.a * g.v + f.v * g.a); }
__device__ __forceinline__ Dual operator/(const Dual& f, const Dual& g) {
  const double gi = 1.0 / g.a, q = f.a * gi; return Dual(q, (f.v - q * g.v) * gi); }
__device__ __forceinline__ Dual operator+(const Dual& f, double s) { return Dual(f.a + s, f.v); }
__device__ __forceinline__ Dual operator+(double s, const Dual& f) { return Dual(f.a + s, f.v); }
__device__ __forceinline__ Dual operator-(const Dual& f, double s) { return Dual(f.a - s, f.v); }
__device__ __forceinline__ Dual operator-(double s, const Dual& f) { return Dual(s - f.a, -f.v); }
__device__ __forceinline__ Dual operator*(const Dual& f, double s) { return Dual(f.a * s, f.v * s); }
__device__ __forceinline__ Dual operator*(double s, const Dual& f) { return Dual(f.a * s, f.v * s); }
__device__ __forceinline__ Dual operator/(const Dual& f, double s) { const double si = 1.0 / s; return Dual(f.a * si, f.v * si); }
__device__ __forceinline__ Dual operator/(double s, const Dual& g) { return Dual(s / g.a, g.v * (-s / (g.a * g.a))); }
__device__ __forceinline__ bool operator>(const Dual& f, double s) { return f.a > s; }
__device__ __forceinline__ bool operator<(const Dual& f, double s) { return f.a < s; }
__device__ __forceinline__ Dual dsqrt(const Dual& f) { const double t = sqrt(f.a); return Dual(t, f.v * (1.0 / (2.0 * t))); }
__device__ __forceinline__ Dual dsin(const Dual& f) { return Dual(sin(f.a), cos(f.a) * f.v); }
__device__ __forceinline__ Dual dcos(const Dual& f) { return Dual(cos(f.a), -sin(f.a) * f.v); }
__device__ __forceinline__ Dual datan2(const Dual& g, const Dual& f) {
  const double t = 1.0 / (f.a * f.a + g.a * g.a); return Dual(atan2(g.a, f.a), t * (-g.a * f.v + f.a * g.v)); }
__device__ __forceinline__ double dsqrt(double x) { return sqrt(x); }
__device__ __forceinline__ double dsin(double x) { return sin(x); }
__device__ __forceinline__ double dcos(double x) { return cos(x); }
__device__ __forceinline__ double datan2(double y, double x) { return atan2(y, x); }

// ---- ceres/rotation.h (1.14 / 2.0), (w,x,y,z) ----
template <typename T> __device__ __forceinline__ void AngleAxisToQuaternion(const T* aa, T* q) {
  const T th2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
  if (th2 > 0.0) {
    const T th = dsqrt(th2); const T half = th * 0.5; const T k = dsin(half) / th;
    q[0] = dcos(half); q[1] = aa[0] * k; q[2] = aa[1] * k; q[3] = aa[2] * k;
  } else { q[0] = T(1.0); q[1] = aa[0] * 0.5; q[2] = aa[1] * 0.5; q[3] = aa[2] * 0.5; }
}
template <typename T> __device__ __forceinline__ void QuaternionToAngleAxis(const T* q, T* aa) {
  const T s2 = q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  if (s2 > 0.0) {
    const T s = dsqrt(s2);
    const T two_theta = 2.0 * ((q[0] < 0.0) ? datan2(-s, -q[0]) : datan2(s, q[0]));
    const T k = two_theta / s;
    aa[0] = q[1] * k; aa[1] = q[2] * k; aa[2] = q[3] * k;
  } else { aa[0] = q[1] * 2.0; aa[1] = q[2] * 2.0; aa[2] = q[3] * 2.0; }
}
template <typename T> __device__ __forceinline__ void QuaternionProduct(const T* z, const T* w, T* zw) {
  zw[0] = z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3];
  zw[1] = z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2];
  zw[2] = z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1];
  zw[3] = z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0];
}
template <typename T> __device__ __forceinline__ void UnitQuaternionRotatePoint(const T* q, const T* pt, T* r) {
  const T t2 = q[0] * q[1], t3 = q[0] * q[2], t4 = q[0] * q[3], t5 = -q[1] * q[1], t6 = q[1] * q[2];
  const T t7 = q[1] * q[3], t8 = -q[2] * q[2], t9 = q[2] * q[3], t1 = -q[3] * q[3];
  r[0] = 2.0 * ((t8 + t1) * pt[0] + (t6 - t4) * pt[1] + (t3 + t7) * pt[2]) + pt[0];
  r[1] = 2.0 * ((t4 + t6) * pt[0] + (t5 + t1) * pt[1] + (t9 - t2) * pt[2]) + pt[1];
  r[2] = 2.0 * ((t7 - t3) * pt[0] + (t2 + t9) * pt[1] + (t5 + t8) * pt[2]) + pt[2];
}
template <typename T> __device__ __forceinline__ void QuaternionRotatePoint(const T* q, const T* pt, T* r) {
  const T scale = 1.0 / dsqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const T unit[4] = {scale * q[0], scale * q[1], scale * q[2], scale * q[3]};
  UnitQuaternionRotatePoint(unit, pt, r);
}
// ---- Eigen::Quaternion<T> generic-scalar operations used by the IMU functor ----
template <typename T> __device__ __forceinline__ void EigenQuatProduct(const T* a, const T* b, T* r) {
  r[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  r[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  r[2] = a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3];
  r[3] = a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1];
}
template <typename T> __device__ __forceinline__ void EigenQuatInverse(const T* q, T* r) {
  const T n2 = (q[1] * q[1] + q[2] * q[2]) + (q[3] * q[3] + q[0] * q[0]);
  if (n2 > 0.0) { r[0] = q[0] / n2; r[1] = -q[1] / n2; r[2] = -q[2] / n2; r[3] = -q[3] / n2; }
  else { r[0] = T(0.0); r[1] = T(0.0); r[2] = T(0.0); r[3] = T(0.0); }
}
template <typename T> __device__ __forceinline__ void EigenQuatRotate(const T* q, const T* v, T* r) {
  T uv[3] = {q[2] * v[2] - q[3] * v[1], q[3] * v[0] - q[1] * v[2], q[1] * v[1] - q[2] * v[0]};
  uv[0] = uv[0] + uv[0]; uv[1] = uv[1] + uv[1]; uv[2] = uv[2] + uv[2];
  r[0] = v[0] + q[0] * uv[0] + (q[2] * uv[2] - q[3] * uv[1]);
  r[1] = v[1] + q[0] * uv[1] + (q[3] * uv[0] - q[1] * uv[2]);
  r[2] = v[2] + q[0] * uv[2] + (q[1] * uv[1] - q[2] * uv[0]);
}
template <typename T> __device__ __forceinline__ void Mat3Vec(const double* M, const T* v, T* r) {
#pragma unroll
  for (int i = 0; i < 3; ++i) r[i] = M[3 * i] * v[0] + (M[3 * i + 1] * v[1] + M[3 * i + 2] * v[2]);
}
template <typename T, int NR> __device__ __forceinline__ void ApplyLeft(const double* __restrict__ A, T* r) {
  T tmp[NR];
  for (int i = 0; i < NR; ++i) {
    T acc = A[i * NR] * r[0];
    for (int j = 1; j < NR; ++j) acc = acc + A[i * NR + j] * r[j];
    tmp[i] = acc;
  }
  for (int i = 0; i < NR; ++i) r[i] = tmp[i];
}

// ---- cost functors (same block / constant layouts as include/bsnls.h) ----
// fuse_constraints/normal_prior_orientation_3d_cost_functor.h:93-122 (A applied by caller)
template <typename T> __device__ __forceinline__ void PriorOrientationDelta(const T* q, const double* b, T* res3) {
  T obs_inv[4] = {T(b[0]), T(-b[1]), T(-b[2]), T(-b[3])}; T diff[4];
  QuaternionProduct(obs_inv, q, diff); QuaternionToAngleAxis(diff, res3);
}
// fuse_constraints/normal_delta_orientation_3d_cost_functor.h:92-123
template <typename T> __device__ __forceinline__ void DeltaOrientation(const T* q1, const T* q2, const double* b, T* res3) {
  T q1_inv[4] = {q1[0], -q1[1], -q1[2], -q1[3]}; T obs_inv[4] = {T(b[0]), T(-b[1]), T(-b[2]), T(-b[3])};
  T diff[4], err[4];
  QuaternionProduct(q1_inv, q2, diff); QuaternionProduct(obs_inv, diff, err); QuaternionToAngleAxis(err, res3);
}
// fuse_constraints/normal_delta_pose_3d_cost_functor.h:107-148
template <typename T> __device__ __forceinline__ void DeltaPose(const double* c, const T* p1, const T* q1, const T* p2, const T* q2, T* r) {
  T q1_inv[4] = {q1[0], -q1[1], -q1[2], -q1[3]}; T pd[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]}; T rot[3];
  QuaternionRotatePoint(q1_inv, pd, rot);
  r[0] = rot[0] - c[0]; r[1] = rot[1] - c[1]; r[2] = rot[2] - c[2];
  DeltaOrientation(q1, q2, c + 3, r + 3);
  ApplyLeft<T, 6>(c + 7, r);
}
// bs_constraints/relative_pose/delta_pose_3d_with_extrinsics_cost_functor.h:65-109
template <typename T> __device__ __forceinline__ void ToSensorFrame(const T* pWB, const T* oWB, const T* pBS, const T* oBS, T* pWS, T* oWS) {
  QuaternionProduct(oWB, oBS, oWS); T t[3]; QuaternionRotatePoint(oWB, pBS, t);
  pWS[0] = t[0] + pWB[0]; pWS[1] = t[1] + pWB[1]; pWS[2] = t[2] + pWB[2];
}
// bs_constraints/inertial/normal_delta_imu_state_3d_cost_functor.h:59-141; p = 32 ambient parameters in block order
template <typename T> __device__ __forceinline__ void ImuDelta(const double* __restrict__ c, const T* p, T* r) {
  const T* q_i = p; const T* p_i = p + 4; const T* v_i = p + 7; const T* bg_i = p + 10; const T* ba_i = p + 13;
  const T* q_j = p + 16; const T* p_j = p + 20; const T* v_j = p + 23; const T* bg_j = p + 26; const T* ba_j = p + 29;
  const double dt = c[0];
  const double G[3] = {0.0, 0.0, -9.80665};
  T dbg[3], dba[3];
  for (int k = 0; k < 3; ++k) { dbg[k] = bg_i[k] - c[56 + k]; dba[k] = ba_i[k] - c[59 + k]; }
  T q_tmp[3]; Mat3Vec(c + 11, dbg, q_tmp);
  T dQ[4] = {T(1.0), q_tmp[0] / 2.0, q_tmp[1] / 2.0, q_tmp[2] / 2.0};
  T dqT[4] = {T(c[1]), T(c[2]), T(c[3]), T(c[4])};
  T q_corr[4]; EigenQuatProduct(dqT, dQ, q_corr);
  T t1[3], t2[3], p_corr[3], v_corr[3];
  Mat3Vec(c + 20, dbg, t1); Mat3Vec(c + 29, dba, t2);
  for (int k = 0; k < 3; ++k) p_corr[k] = (c[5 + k] + t1[k]) + t2[k];
  Mat3Vec(c + 38, dbg, t1); Mat3Vec(c + 47, dba, t2);
  for (int k = 0; k < 3; ++k) v_corr[k] = (c[8 + k] + t1[k]) + t2[k];
  T qi_inv[4], qc_inv[4], qij[4], qe[4];
  EigenQuatInverse(q_i, qi_inv); EigenQuatProduct(qi_inv, q_j, qij);
  EigenQuatInverse(q_corr, qc_inv); EigenQuatProduct(qc_inv, qij, qe);
  r[0] = 2.0 * qe[1]; r[1] = 2.0 * qe[2]; r[2] = 2.0 * qe[3];
  T qi_conj[4] = {q_i[0], -q_i[1], -q_i[2], -q_i[3]}; T arg[3], rot[3];
  for (int k = 0; k < 3; ++k) arg[k] = ((p_j[k] - p_i[k]) - dt * v_i[k]) - ((0.5 * dt) * dt) * G[k];
  EigenQuatRotate(qi_conj, arg, rot);
  for (int k = 0; k < 3; ++k) r[3 + k] = rot[k] - p_corr[k];
  for (int k = 0; k < 3; ++k) arg[k] = (v_j[k] - v_i[k]) - dt * G[k];
  EigenQuatRotate(qi_conj, arg, rot);
  for (int k = 0; k < 3; ++k) r[6 + k] = rot[k] - v_corr[k];
  for (int k = 0; k < 3; ++k) { r[9 + k] = bg_j[k] - bg_i[k]; r[12 + k] = ba_j[k] - ba_i[k]; }
  ApplyLeft<T, 15>(c + 62, r);
}
// fuse_constraints/src/marginal_cost_function.cpp:58-113: r = b + sum_i A_i Minus(x_bar_i, x_i). The reference's Jacobian
// is A_i times the Minus Jacobian evaluated at x_i (orientation_3d_stamped.h:130-142) -- not the derivative of the angle-axis
// difference -- so the dual part of a quaternion block's delta is that linear map applied to the dual parts of x.
__device__ __forceinline__ double TVal(double x) { return x; }
__device__ __forceinline__ double TVal(const Dual& x) { return x.a; }
__device__ __forceinline__ double TDer(double) { return 0.0; }
__device__ __forceinline__ double TDer(const Dual& x) { return x.v; }
__device__ __forceinline__ void TSet(double& t, double a, double) { t = a; }
__device__ __forceinline__ void TSet(Dual& t, double a, double v) { t = Dual(a, v); }
template <typename T> __device__ __forceinline__ void MarginalFactor(const double* __restrict__ c, const T* p, T* r) {
  const int nb = (int)c[0];
  const double* bvec = c + 11; const double* xbar = c + 26; const double* A = c + 58;
  for (int i = 0; i < BSNLS_MARGINAL_MAX_ROWS; ++i) r[i] = T(bvec[i]);
  int off = 0;
  for (int b = 0; b < nb; ++b) {
    const int sz = (int)c[1 + b];
    T d[3];
    if (sz == 4) {
      const double x[4] = {TVal(p[off]), TVal(p[off + 1]), TVal(p[off + 2]), TVal(p[off + 3])};
      const double dx[4] = {TDer(p[off]), TDer(p[off + 1]), TDer(p[off + 2]), TDer(p[off + 3])};
      double aa[3]; PriorOrientationDelta<double>(x, xbar + off, aa);
      TSet(d[0], aa[0], 2.0 * (-x[1] * dx[0] + x[0] * dx[1] + x[3] * dx[2] - x[2] * dx[3]));
      TSet(d[1], aa[1], 2.0 * (-x[2] * dx[0] - x[3] * dx[1] + x[0] * dx[2] + x[1] * dx[3]));
      TSet(d[2], aa[2], 2.0 * (-x[3] * dx[0] + x[2] * dx[1] - x[1] * dx[2] + x[0] * dx[3]));
    } else {
      for (int k = 0; k < 3; ++k) d[k] = p[off + k] - xbar[off + k];
    }
    for (int i = 0; i < BSNLS_MARGINAL_MAX_ROWS; ++i) {
      const double* a = A + i * 30 + 3 * b;
      r[i] = r[i] + (a[0] * d[0] + a[1] * d[1] + a[2] * d[2]);
    }
    off += sz;
  }
}
// Dispatch over the non-reprojection factor types (ids of bsnls_factor_type); p = concatenated ambient blocks.
template <typename T> __device__ __forceinline__ void EvalGenericFactor(int type, const double* __restrict__ c, const T* p, T* r) {
  switch (type) {
    case 1: ImuDelta(c, p, r); break;
    case 2: {  // NormalPriorImuState3DCostFunctor (normal_prior_imu_state_3d_cost_functor.h:54-90)
      PriorOrientationDelta(p, c, r);
      for (int k = 0; k < 12; ++k) r[3 + k] = p[4 + k] - c[4 + k];
      ApplyLeft<T, 15>(c + 16, r); break; }
    case 3: DeltaPose(c, p, p + 3, p + 7, p + 10, r); break;
    case 4: {
      T p1[3], o1[4], p2[3], o2[4];
      ToSensorFrame(p, p + 3, p + 14, p + 17, p1, o1); ToSensorFrame(p + 7, p + 10, p + 14, p + 17, p2, o2);
      DeltaPose(c, p1, o1, p2, o2, r); break; }
    case 5: {  // NormalPriorPose3DCostFunctor (normal_prior_pose_3d_cost_functor.h:86-123)
      r[0] = p[0] - c[0]; r[1] = p[1] - c[1]; r[2] = p[2] - c[2];
      PriorOrientationDelta(p + 3, c + 3, r + 3); ApplyLeft<T, 6>(c + 7, r); break; }
    case 6: DeltaOrientation(p, p + 4, c, r); ApplyLeft<T, 3>(c + 4, r); break;
    case 7: for (int k = 0; k < 3; ++k) r[k] = p[k] - c[k]; ApplyLeft<T, 3>(c + 3, r); break;
    case 8: for (int k = 0; k < 3; ++k) r[k] = (p[3 + k] - p[k]) - c[k]; ApplyLeft<T, 3>(c + 3, r); break;
    case 9: MarginalFactor(c, p, r); break;
    case 10: {  // GravityAlignmentCostFunctor (gravity_alignment_cost_functor.h:53-66)
      const T g[3] = {T(c[0]), T(c[1]), T(c[2])}; T w[3]; QuaternionRotatePoint(p, g, w);
      r[0] = c[3] * w[0] + c[4] * w[1]; r[1] = c[5] * w[0] + c[6] * w[1]; break; }
    case 11: PriorOrientationDelta(p, c, r); ApplyLeft<T, 3>(c + 4, r); break;   // NormalPriorOrientation3DCostFunctor
    default: break;
  }
}

}  // namespace bsnls
