// bs_common::PreIntegrator restated for the host side: it produces the constants of the IMU factor
// (bs_common/src/bs_common/preintegrator.cpp:26-144, include/bs_common/preintegrator.h:13-143).
// Runs once per factor creation, never inside the solve.
#pragma once
#include <vector>

namespace bsfuse {

struct IMUData { double t; double w[3]; double a[3]; };

class PreIntegrator {
 public:
  // error-state order (preintegrator.h:13-20)
  enum { ES_Q = 0, ES_P = 3, ES_V = 6, ES_BG = 9, ES_BA = 12, ES_SIZE = 15 };
  double cov_tol = 1e-5, bias_cov_tol = 1e-9, invalid_inv_cov_weight = 1e-4;
  double cov_w = 0.00057, cov_a = 0.00094, cov_bg = 0.0000037, cov_ba = 0.0000024;  // x I3 (ig2/imu.json)
  struct Delta { double t; double q[4]; double p[3]; double v[3]; double cov[225]; double sqrt_inv_cov[225]; } delta;
  struct Jacobian { double dq_dbg[9], dp_dbg[9], dp_dba[9], dv_dbg[9], dv_dba[9]; } jacobian;
  std::vector<IMUData> data;  // sorted by time (std::map<ros::Time, IMUData> in the reference)

  void Reset();
  void Increment(double dt, const IMUData& d, const double bg[3], const double ba[3], bool compute_jacobian,
                 bool compute_covariance);
  bool Integrate(double t, const double bg[3], const double ba[3], bool compute_jacobian, bool compute_covariance,
                 bool compute_information);
  void ComputeSqrtInvCov();
  // Fill the BSNLS_IMU_DELTA_STRIDE constant block of include/bsnls.h (A = info_weight * sqrt_inv_cov,
  // normal_delta_imu_state_3d_cost_functor.h:57).
  void FillFactorConstants(const double bg0[3], const double ba0[3], double info_weight, double out[288]) const;
};

}  // namespace bsfuse
