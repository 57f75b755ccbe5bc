// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing under beam_slam_b200/ may include, link or call this.
//
// CPU restatement of the arithmetic the reference hands to Ceres for the fuse-graph solve. The algorithm of
// the solve itself lives in Ceres Solver, which is NOT vendored in /root/reference (find_package(Ceres),
// bs_constraints/CMakeLists.txt:31; un-pinned, API implies 1.13 <= version <= 2.1). What is restated here from
// Ceres' published sources (1.14 / 2.0): ceres/jet.h dual numbers, ceres/rotation.h helpers,
// residual_block.cc + corrector.cc (also restated in-tree by fuse:
// dependencies/fuse/fuse_constraints/src/marginalize_variables.cpp:281-399), trust_region_minimizer.cc,
// levenberg_marquardt_strategy.cc, schur_eliminator_impl.h.
//
// Parity pinning: factor/manifold/linearisation math is pinned by the reference's golden vectors
// (tests/test_oracle_golden.py); per-iteration LM quantities are NOT pinned by any reference test and no Ceres
// binary exists in this image => iteration-level parity with Ceres is "parity unpinned" (see DESIGN.md).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

// ---------------------------------------------------------------------------------------------------
// ceres::Jet<double, N> (ceres/jet.h): value + N partial derivatives, forward mode.
// ---------------------------------------------------------------------------------------------------
template <int N>
struct Jet {
  double a;
  double v[N];
  Jet() : a(0.0) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
  Jet(double s) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; }  // NOLINT
  Jet(double s, int k) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a + g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a - g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f) {
  Jet<N> h; h.a = -f.a; for (int i = 0; i < N; ++i) h.v[i] = -f.v[i]; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a * g.a; for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
  const double g_a_inverse = 1.0 / g.a;
  const double f_a_by_g_a = f.a * g_a_inverse;
  Jet<N> h; h.a = f_a_by_g_a;
  for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - f_a_by_g_a * g.v[i]) * g_a_inverse;
  return h; }
template <int N> inline Jet<N> operator+(const Jet<N>& f, double s) { Jet<N> h = f; h.a += s; return h; }
template <int N> inline Jet<N> operator+(double s, const Jet<N>& f) { Jet<N> h = f; h.a += s; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, double s) { Jet<N> h = f; h.a -= s; return h; }
template <int N> inline Jet<N> operator-(double s, const Jet<N>& f) {
  Jet<N> h; h.a = s - f.a; for (int i = 0; i < N; ++i) h.v[i] = -f.v[i]; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, double s) {
  Jet<N> h; h.a = f.a * s; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s; return h; }
template <int N> inline Jet<N> operator*(double s, const Jet<N>& f) { return f * s; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, double s) {
  const double s_inverse = 1.0 / s; return f * s_inverse; }
template <int N> inline Jet<N> operator/(double s, const Jet<N>& g) {
  const double minus_s_g_a_inverse2 = -s / (g.a * g.a);
  Jet<N> h; h.a = s / g.a; for (int i = 0; i < N; ++i) h.v[i] = g.v[i] * minus_s_g_a_inverse2; return h; }
template <int N> inline Jet<N>& operator+=(Jet<N>& f, const Jet<N>& g) { f = f + g; return f; }
template <int N> inline Jet<N>& operator-=(Jet<N>& f, const Jet<N>& g) { f = f - g; return f; }
template <int N> inline bool operator>(const Jet<N>& f, double s) { return f.a > s; }
template <int N> inline bool operator<(const Jet<N>& f, double s) { return f.a < s; }
template <int N> inline Jet<N> sqrt(const Jet<N>& f) {
  const double tmp = std::sqrt(f.a); const double two_a_inverse = 1.0 / (2.0 * tmp);
  Jet<N> h; h.a = tmp; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * two_a_inverse; return h; }
template <int N> inline Jet<N> sin(const Jet<N>& f) {
  const double c = std::cos(f.a); Jet<N> h; h.a = std::sin(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = c * f.v[i]; return h; }
template <int N> inline Jet<N> cos(const Jet<N>& f) {
  const double ms = -std::sin(f.a); Jet<N> h; h.a = std::cos(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = ms * f.v[i]; return h; }
template <int N> inline Jet<N> atan2(const Jet<N>& g, const Jet<N>& f) {
  const double tmp = 1.0 / (f.a * f.a + g.a * g.a);
  Jet<N> h; h.a = std::atan2(g.a, f.a);
  for (int i = 0; i < N; ++i) h.v[i] = tmp * (-g.a * f.v[i] + f.a * g.v[i]);
  return h; }
inline double sqrt(double x) { return std::sqrt(x); }
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }
inline double atan2(double y, double x) { return std::atan2(y, x); }

// ---------------------------------------------------------------------------------------------------
// ceres/rotation.h (1.14 / 2.0). Quaternions are (w,x,y,z).
// ---------------------------------------------------------------------------------------------------
template <typename T> inline void AngleAxisToQuaternion(const T* aa, T* q) {
  const T& a0 = aa[0]; const T& a1 = aa[1]; const T& a2 = aa[2];
  const T theta_squared = a0 * a0 + a1 * a1 + a2 * a2;
  if (theta_squared > 0.0) {
    const T theta = sqrt(theta_squared);
    const T half_theta = theta * 0.5;
    const T k = sin(half_theta) / theta;
    q[0] = cos(half_theta); q[1] = a0 * k; q[2] = a1 * k; q[3] = a2 * k;
  } else {
    const double k = 0.5;
    q[0] = T(1.0); q[1] = a0 * k; q[2] = a1 * k; q[3] = a2 * k;
  }
}
template <typename T> inline void QuaternionToAngleAxis(const T* q, T* aa) {
  const T& q1 = q[1]; const T& q2 = q[2]; const T& q3 = q[3];
  const T sin_squared_theta = q1 * q1 + q2 * q2 + q3 * q3;
  if (sin_squared_theta > 0.0) {
    const T sin_theta = sqrt(sin_squared_theta);
    const T& cos_theta = q[0];
    const T two_theta = 2.0 * ((cos_theta < 0.0) ? atan2(-sin_theta, -cos_theta)
                                                 : atan2(sin_theta, cos_theta));
    const T k = two_theta / sin_theta;
    aa[0] = q1 * k; aa[1] = q2 * k; aa[2] = q3 * k;
  } else {
    const double k = 2.0;
    aa[0] = q1 * k; aa[1] = q2 * k; aa[2] = q3 * k;
  }
}
template <typename T> inline void QuaternionProduct(const T* z, const T* w, T* zw) {
  zw[0] = z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3];
  zw[1] = z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2];
  zw[2] = z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1];
  zw[3] = z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0];
}
template <typename T> inline void UnitQuaternionRotatePoint(const T* q, const T* pt, T* result) {
  const T t2 = q[0] * q[1]; const T t3 = q[0] * q[2]; const T t4 = q[0] * q[3];
  const T t5 = -q[1] * q[1]; const T t6 = q[1] * q[2]; const T t7 = q[1] * q[3];
  const T t8 = -q[2] * q[2]; const T t9 = q[2] * q[3]; const T t1 = -q[3] * q[3];
  result[0] = 2.0 * ((t8 + t1) * pt[0] + (t6 - t4) * pt[1] + (t3 + t7) * pt[2]) + pt[0];
  result[1] = 2.0 * ((t4 + t6) * pt[0] + (t5 + t1) * pt[1] + (t9 - t2) * pt[2]) + pt[1];
  result[2] = 2.0 * ((t7 - t3) * pt[0] + (t2 + t9) * pt[1] + (t5 + t8) * pt[2]) + pt[2];
}
template <typename T> inline void QuaternionRotatePoint(const T* q, const T* pt, T* result) {
  const T scale = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const T unit[4] = {scale * q[0], scale * q[1], scale * q[2], scale * q[3]};
  UnitQuaternionRotatePoint(unit, pt, result);
}

// ---------------------------------------------------------------------------------------------------
// Eigen::Quaternion<T> operations used by the IMU functor (Eigen/src/Geometry/Quaternion.h, generic
// scalar path): product, conjugate, inverse (= conjugate / squaredNorm), q * v (_transformVector).
// Stored here as (w,x,y,z).
// ---------------------------------------------------------------------------------------------------
template <typename T> inline void EigenQuatProduct(const T* a, const T* b, T* r) {
  r[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  r[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  r[2] = a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3];
  r[3] = a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1];
}
template <typename T> inline void EigenQuatInverse(const T* q, T* r) {
  const T n2 = (q[1] * q[1] + q[2] * q[2]) + (q[3] * q[3] + q[0] * q[0]);
  if (n2 > 0.0) { r[0] = q[0] / n2; r[1] = -q[1] / n2; r[2] = -q[2] / n2; r[3] = -q[3] / n2; }
  else { r[0] = T(0.0); r[1] = T(0.0); r[2] = T(0.0); r[3] = T(0.0); }
}
template <typename T> inline void EigenQuatRotate(const T* q, const T* v, T* r) {
  // uv = q.vec x v; uv += uv; r = v + w*uv + q.vec x uv
  T uv[3] = {q[2] * v[2] - q[3] * v[1], q[3] * v[0] - q[1] * v[2], q[1] * v[1] - q[2] * v[0]};
  uv[0] = uv[0] + uv[0]; uv[1] = uv[1] + uv[1]; uv[2] = uv[2] + uv[2];
  r[0] = v[0] + q[0] * uv[0] + (q[2] * uv[2] - q[3] * uv[1]);
  r[1] = v[1] + q[0] * uv[1] + (q[3] * uv[0] - q[1] * uv[2]);
  r[2] = v[2] + q[0] * uv[2] + (q[1] * uv[1] - q[2] * uv[0]);
}
// Eigen fixed-size 3x3 (row-major storage here) times 3-vector: coefficient-based lazy product,
// 3-term sums associate as a0 + (a1 + a2) (Eigen/src/Core/Redux.h redux_novec_unroller).
template <typename T> inline void Mat3Vec(const double* M, const T* v, T* r) {
  for (int i = 0; i < 3; ++i) r[i] = M[3 * i] * v[0] + (M[3 * i + 1] * v[1] + M[3 * i + 2] * v[2]);
}

}  // namespace orc
