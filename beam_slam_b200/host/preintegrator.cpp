#include "preintegrator.h"

#include <algorithm>
#include <cmath>
#include <cstring>

#include "linalg.h"

namespace bsfuse {
using namespace linalg;

void PreIntegrator::Reset() {
  std::memset(&delta, 0, sizeof(delta)); delta.q[0] = 1.0;
  std::memset(&jacobian, 0, sizeof(jacobian));
}

static void AddScaled(double* dst, const double* src, double s, int n) { for (int i = 0; i < n; ++i) dst[i] += s * src[i]; }
static void SetBlock(double* M, int ld, int r, int c, const double* B, double s) {
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) M[(r + i) * ld + c + j] = s * B[i * 3 + j];
}

// preintegrator.cpp:26-96
void PreIntegrator::Increment(double dtd, const IMUData& d, const double bg[3], const double ba[3],
                              bool compute_jacobian, bool compute_covariance) {
  const double w[3] = {d.w[0] - bg[0], d.w[1] - bg[1], d.w[2] - bg[2]};
  const double a[3] = {d.a[0] - ba[0], d.a[1] - ba[1], d.a[2] - ba[2]};
  const double wdt[3] = {w[0] * dtd, w[1] * dtd, w[2] * dtd};
  const double whalf[3] = {0.5 * w[0] * dtd, 0.5 * w[1] * dtd, 0.5 * w[2] * dtd};
  double Rfull[9], Rhalf[9], RfullT[9], Rd[9], Sa[9], RdSa[9], Jr[9];
  ExpSO3(wdt, Rfull); ExpSO3(whalf, Rhalf); Transpose(Rfull, 3, 3, RfullT);
  QuatToRot(delta.q, Rd); Skew(a, Sa); MatMul(Rd, Sa, 3, 3, 3, RdSa); RightJacobianSO3(wdt, Jr);
  const double I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};

  if (compute_covariance) {
    double A[81] = {0}, B[54] = {0};
    for (int i = 0; i < 9; ++i) A[i * 9 + i] = 1.0;
    SetBlock(A, 9, ES_Q, ES_Q, RfullT, 1.0);
    SetBlock(A, 9, ES_V, ES_Q, RdSa, -dtd);
    SetBlock(A, 9, ES_P, ES_Q, RdSa, -0.5 * dtd * dtd);
    SetBlock(A, 9, ES_P, ES_V, I3, dtd);
    SetBlock(B, 6, ES_Q, 0, Jr, dtd);
    SetBlock(B, 6, ES_V, 3, Rd, dtd);
    SetBlock(B, 6, ES_P, 3, Rd, 0.5 * dtd * dtd);
    const double inv_dtd = 1.0 / std::max(dtd, 1.0e-7);
    double W[36] = {0};
    for (int i = 0; i < 3; ++i) { W[i * 6 + i] = cov_w * inv_dtd; W[(3 + i) * 6 + 3 + i] = cov_a * inv_dtd; }
    double C9[81], AC[81], AT[81], ACA[81], BW[54], BT[54], BWB[81];
    for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) C9[i * 9 + j] = delta.cov[i * 15 + j];
    MatMul(A, C9, 9, 9, 9, AC); Transpose(A, 9, 9, AT); MatMul(AC, AT, 9, 9, 9, ACA);
    MatMul(B, W, 9, 6, 6, BW); Transpose(B, 9, 6, BT); MatMul(BW, BT, 9, 6, 9, BWB);
    for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) delta.cov[i * 15 + j] = ACA[i * 9 + j] + BWB[i * 9 + j];
    for (int i = 0; i < 3; ++i) { delta.cov[(ES_BG + i) * 15 + ES_BG + i] += cov_bg * dtd; delta.cov[(ES_BA + i) * 15 + ES_BA + i] += cov_ba * dtd; }
  }
  if (compute_jacobian) {
    double T[9];
    // dp_dbg += dt*dv_dbg - 0.5 dt^2 Rd Sa dq_dbg
    MatMul(RdSa, jacobian.dq_dbg, 3, 3, 3, T);
    AddScaled(jacobian.dp_dbg, jacobian.dv_dbg, dtd, 9); AddScaled(jacobian.dp_dbg, T, -0.5 * dtd * dtd, 9);
    // dp_dba += dt*dv_dba - 0.5 dt^2 Rd
    AddScaled(jacobian.dp_dba, jacobian.dv_dba, dtd, 9); AddScaled(jacobian.dp_dba, Rd, -0.5 * dtd * dtd, 9);
    // dv_dbg -= dt Rd Sa dq_dbg ; dv_dba -= dt Rd
    AddScaled(jacobian.dv_dbg, T, -dtd, 9);
    AddScaled(jacobian.dv_dba, Rd, -dtd, 9);
    // dq_dbg = Rfull^T dq_dbg - dt Jr
    MatMul(RfullT, jacobian.dq_dbg, 3, 3, 3, T);
    for (int i = 0; i < 9; ++i) jacobian.dq_dbg[i] = T[i] - dtd * Jr[i];
  }
  double Rmid[9], amid[3];
  MatMul(Rd, Rhalf, 3, 3, 3, Rmid);
  for (int i = 0; i < 3; ++i) amid[i] = Rmid[3 * i] * a[0] + Rmid[3 * i + 1] * a[1] + Rmid[3 * i + 2] * a[2];
  delta.t += dtd;
  for (int i = 0; i < 3; ++i) { delta.p[i] = delta.p[i] + dtd * delta.v[i] + 0.5 * dtd * dtd * amid[i]; }
  for (int i = 0; i < 3; ++i) delta.v[i] = delta.v[i] + dtd * amid[i];
  double qfull[4], qn[4]; RotToQuat(Rfull, qfull); QuatMul(delta.q, qfull, qn);
  const double n = std::sqrt(qn[0] * qn[0] + qn[1] * qn[1] + qn[2] * qn[2] + qn[3] * qn[3]);
  for (int i = 0; i < 4; ++i) delta.q[i] = qn[i] / n;
}

// preintegrator.cpp:98-122
bool PreIntegrator::Integrate(double t, const double bg[3], const double ba[3], bool cj, bool cc, bool ci) {
  if (data.empty()) return false;
  Reset();
  for (size_t i = 0; i + 1 < data.size(); ++i) {
    if (data[i + 1].t > t) break;
    Increment(data[i + 1].t - data[i].t, data[i], bg, ba, cj, cc);
  }
  const double dt = t - data.back().t;
  if (dt > 0) Increment(dt, data.back(), bg, ba, cj, cc);
  if (ci) ComputeSqrtInvCov();
  return true;
}

// preintegrator.cpp:124-144
void PreIntegrator::ComputeSqrtInvCov() {
  double n1 = 0, n2 = 0;
  for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) n1 += delta.cov[i * 15 + j] * delta.cov[i * 15 + j];
  if (std::sqrt(n1) < cov_tol) for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) delta.cov[i * 15 + j] = (i == j) ? cov_tol : 0.0;
  for (int i = 9; i < 15; ++i) for (int j = 9; j < 15; ++j) n2 += delta.cov[i * 15 + j] * delta.cov[i * 15 + j];
  if (std::sqrt(n2) < bias_cov_tol) for (int i = 9; i < 15; ++i) for (int j = 9; j < 15; ++j) delta.cov[i * 15 + j] = (i == j) ? bias_cov_tol : 0.0;
  double inv[225], L[225];
  bool ok = Inverse(delta.cov, 15, inv) && CholeskyLower(inv, 15, L);
  if (ok) {
    for (int i = 0; i < 15; ++i) for (int j = 0; j < 15; ++j) { delta.sqrt_inv_cov[i * 15 + j] = L[j * 15 + i]; ok = ok && std::isfinite(L[j * 15 + i]); }
  }
  if (!ok) for (int i = 0; i < 15; ++i) for (int j = 0; j < 15; ++j) delta.sqrt_inv_cov[i * 15 + j] = (i == j) ? invalid_inv_cov_weight : 0.0;
}

void PreIntegrator::FillFactorConstants(const double bg0[3], const double ba0[3], double w, double out[288]) const {
  std::memset(out, 0, sizeof(double) * 288);
  out[0] = delta.t;
  std::memcpy(out + 1, delta.q, 32); std::memcpy(out + 5, delta.p, 24); std::memcpy(out + 8, delta.v, 24);
  std::memcpy(out + 11, jacobian.dq_dbg, 72); std::memcpy(out + 20, jacobian.dp_dbg, 72);
  std::memcpy(out + 29, jacobian.dp_dba, 72); std::memcpy(out + 38, jacobian.dv_dbg, 72);
  std::memcpy(out + 47, jacobian.dv_dba, 72);
  std::memcpy(out + 56, bg0, 24); std::memcpy(out + 59, ba0, 24);
  for (int i = 0; i < 225; ++i) out[62 + i] = w * delta.sqrt_inv_cov[i];
}

}  // namespace bsfuse
