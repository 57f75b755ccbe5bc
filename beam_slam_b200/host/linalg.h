// Small dense row-major helpers for the host side (constraint constructors, IMU preintegration, generator).
#pragma once
#include <cmath>
#include <vector>

namespace bsfuse {
namespace linalg {

// C[m x n] = A[m x k] * B[k x n]
inline void MatMul(const double* A, const double* B, int m, int k, int n, double* C) {
  for (int i = 0; i < m; ++i)
    for (int j = 0; j < n; ++j) {
      double s = 0.0;
      for (int p = 0; p < k; ++p) s += A[i * k + p] * B[p * n + j];
      C[i * n + j] = s;
    }
}
inline void Transpose(const double* A, int m, int n, double* At) {
  for (int i = 0; i < m; ++i) for (int j = 0; j < n; ++j) At[j * m + i] = A[i * n + j];
}
// Gauss-Jordan with partial pivoting
inline bool Inverse(const double* A, int n, double* inv) {
  std::vector<double> M(A, A + n * n);
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) inv[i * n + j] = (i == j) ? 1.0 : 0.0;
  for (int c = 0; c < n; ++c) {
    int piv = c; double best = std::fabs(M[c * n + c]);
    for (int r = c + 1; r < n; ++r) if (std::fabs(M[r * n + c]) > best) { best = std::fabs(M[r * n + c]); piv = r; }
    if (!(best > 0.0) || !std::isfinite(best)) return false;
    if (piv != c) for (int j = 0; j < n; ++j) { std::swap(M[c * n + j], M[piv * n + j]); std::swap(inv[c * n + j], inv[piv * n + j]); }
    const double d = 1.0 / M[c * n + c];
    for (int j = 0; j < n; ++j) { M[c * n + j] *= d; inv[c * n + j] *= d; }
    for (int r = 0; r < n; ++r) {
      if (r == c) continue;
      const double f = M[r * n + c];
      if (f == 0.0) continue;
      for (int j = 0; j < n; ++j) { M[r * n + j] -= f * M[c * n + j]; inv[r * n + j] -= f * inv[c * n + j]; }
    }
  }
  return true;
}
inline bool CholeskyLower(const double* A, int n, double* L) {
  for (int i = 0; i < n * n; ++i) L[i] = 0.0;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = A[i * n + j];
      for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
      if (i == j) { if (!(s > 0.0) || !std::isfinite(s)) return false; L[i * n + i] = std::sqrt(s); }
      else L[i * n + j] = s / L[j * n + j];
    }
  return true;
}

// ---- SO(3) helpers (libbeam beam_utils/math.h,se3.h is not vendored; these four are all the path touches:
// bs_constraints/src/jacobians.cpp:178, bs_common/src/bs_common/preintegrator.cpp:35-52) ----
inline void Skew(const double v[3], double S[9]) {
  S[0] = 0; S[1] = -v[2]; S[2] = v[1]; S[3] = v[2]; S[4] = 0; S[5] = -v[0]; S[6] = -v[1]; S[7] = v[0]; S[8] = 0;
}
// beam::LieAlgebraToR: Rodrigues exponential map
inline void ExpSO3(const double w[3], double R[9]) {
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
  double a, b;
  if (th < 1e-10) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; } else { a = std::sin(th) / th; b = (1.0 - std::cos(th)) / th2; }
  double K[9]; Skew(w, K); double K2[9]; MatMul(K, K, 3, 3, 3, K2);
  for (int i = 0; i < 9; ++i) R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + a * K[i] + b * K2[i];
}
// beam::RightJacobianOfSO3
inline void RightJacobianSO3(const double w[3], double J[9]) {
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
  double a, b;
  if (th < 1e-10) { a = 0.5 - th2 / 24.0; b = 1.0 / 6.0 - th2 / 120.0; }
  else { a = (1.0 - std::cos(th)) / th2; b = (th - std::sin(th)) / (th2 * th); }
  double K[9]; Skew(w, K); double K2[9]; MatMul(K, K, 3, 3, 3, K2);
  for (int i = 0; i < 9; ++i) J[i] = ((i % 4 == 0) ? 1.0 : 0.0) - a * K[i] + b * K2[i];
}
inline void QuatToRot(const double q[4] /*w,x,y,z*/, double R[9]) {
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - z * w); R[2] = 2 * (x * z + y * w);
  R[3] = 2 * (x * y + z * w); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - x * w);
  R[6] = 2 * (x * z - y * w); R[7] = 2 * (y * z + x * w); R[8] = 1 - 2 * (x * x + y * y);
}
inline void RotToQuat(const double R[9], double q[4]) {
  const double tr = R[0] + R[4] + R[8];
  if (tr > 0) { const double s = std::sqrt(tr + 1.0) * 2; q[0] = 0.25 * s; q[1] = (R[7] - R[5]) / s; q[2] = (R[2] - R[6]) / s; q[3] = (R[3] - R[1]) / s; }
  else if (R[0] > R[4] && R[0] > R[8]) { const double s = std::sqrt(1.0 + R[0] - R[4] - R[8]) * 2; q[0] = (R[7] - R[5]) / s; q[1] = 0.25 * s; q[2] = (R[1] + R[3]) / s; q[3] = (R[2] + R[6]) / s; }
  else if (R[4] > R[8]) { const double s = std::sqrt(1.0 + R[4] - R[0] - R[8]) * 2; q[0] = (R[2] - R[6]) / s; q[1] = (R[1] + R[3]) / s; q[2] = 0.25 * s; q[3] = (R[5] + R[7]) / s; }
  else { const double s = std::sqrt(1.0 + R[8] - R[0] - R[4]) * 2; q[0] = (R[3] - R[1]) / s; q[1] = (R[2] + R[6]) / s; q[2] = (R[5] + R[7]) / s; q[3] = 0.25 * s; }
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; ++i) q[i] /= n;
  if (q[0] < 0) for (int i = 0; i < 4; ++i) q[i] = -q[i];
}
inline void QuatMul(const double a[4], const double b[4], double r[4]) {
  r[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  r[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  r[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  r[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
}

}  // namespace linalg
}  // namespace bsfuse
