// Synthetic graphs of the five BASELINE.json configs: seeded scene, smooth trajectory,
// 200 Hz IMU stream -> PreIntegrator constants, reprojection / IMU / relative-pose / prior constraints, inserted
// into the hash graph exactly as the sensor models would (bs_models/src/lib/vision/visual_map.cpp:173-219,
// bs_models/src/lib/imu/imu_preintegration.cpp:246-320, bs_constraints/src/relative_pose/
// pose_3d_stamped_transaction.cpp:28-82, bs_optimizers/src/fixed_lag_smoother.cpp:244-268).
// Deviation from the survey's plan: the trajectory is an analytic smooth curve (closed-form position, Euler-angle
// attitude) instead of a basalt quintic B-spline; everything downstream (IMU synthesis, preintegration) is the same.
#include "synthetic.h"

#include <cmath>
#include <cstring>
#include <random>

#include "linalg.h"
#include "preintegrator.h"

namespace bsfuse {
using namespace linalg;

namespace {

struct Rng {
  std::mt19937_64 g;
  explicit Rng(uint64_t s) : g(s) {}
  double uniform(double a, double b) { return a + (b - a) * ((g() >> 11) * (1.0 / 9007199254740992.0)); }
  double normal() {  // Box-Muller, portable
    const double u1 = std::max(uniform(0, 1), 1e-300), u2 = uniform(0, 1);
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
  }
  int randint(int lo, int hi) { return lo + static_cast<int>(g() % static_cast<uint64_t>(hi - lo + 1)); }
};

constexpr double kOmega = 0.05, kRadius = 20.0;
void Position(double t, double p[3]) {
  p[0] = kRadius * std::cos(kOmega * t) + 0.3 * std::sin(0.9 * t + 1.0);
  p[1] = kRadius * std::sin(kOmega * t) + 0.3 * std::sin(1.1 * t + 2.0);
  p[2] = 0.5 * std::sin(0.3 * t);
}
void Velocity(double t, double v[3]) {
  v[0] = -kRadius * kOmega * std::sin(kOmega * t) + 0.27 * std::cos(0.9 * t + 1.0);
  v[1] = kRadius * kOmega * std::cos(kOmega * t) + 0.33 * std::cos(1.1 * t + 2.0);
  v[2] = 0.15 * std::cos(0.3 * t);
}
void Acceleration(double t, double a[3]) {
  a[0] = -kRadius * kOmega * kOmega * std::cos(kOmega * t) - 0.243 * std::sin(0.9 * t + 1.0);
  a[1] = -kRadius * kOmega * kOmega * std::sin(kOmega * t) - 0.363 * std::sin(1.1 * t + 2.0);
  a[2] = -0.045 * std::sin(0.3 * t);
}
void Attitude(double t, double R[9]) {
  const double yaw = kOmega * t + 1.5707963267948966 + 0.2 * std::sin(0.4 * t);
  const double pitch = 0.1 * std::sin(0.5 * t + 1.0), roll = 0.1 * std::sin(0.7 * t);
  const double cy = std::cos(yaw), sy = std::sin(yaw), cp = std::cos(pitch), sp = std::sin(pitch), cr = std::cos(roll), sr = std::sin(roll);
  const double Rz[9] = {cy, -sy, 0, sy, cy, 0, 0, 0, 1}, Ry[9] = {cp, 0, sp, 0, 1, 0, -sp, 0, cp}, Rx[9] = {1, 0, 0, 0, cr, -sr, 0, sr, cr};
  double T[9]; MatMul(Rz, Ry, 3, 3, 3, T); MatMul(T, Rx, 3, 3, 3, R);
}
void LogSO3(const double R[9], double w[3]) {
  const double c = std::min(1.0, std::max(-1.0, 0.5 * (R[0] + R[4] + R[8] - 1.0)));
  const double th = std::acos(c);
  const double k = th < 1e-8 ? 0.5 + th * th / 12.0 : th / (2.0 * std::sin(th));
  w[0] = k * (R[7] - R[5]); w[1] = k * (R[2] - R[6]); w[2] = k * (R[3] - R[1]);
}
void BodyRate(double t, double w[3]) {  // vee(R^T dR/dt), central difference
  const double h = 1e-5;
  double Ra[9], Rb[9], RaT[9], D[9];
  Attitude(t - h, Ra); Attitude(t + h, Rb); Transpose(Ra, 3, 3, RaT); MatMul(RaT, Rb, 3, 3, 3, D);
  LogSO3(D, w);
  for (int i = 0; i < 3; ++i) w[i] /= (2 * h);
}
void MatVec3(const double* M, const double* v, double* r) { for (int i = 0; i < 3; ++i) r[i] = M[3 * i] * v[0] + M[3 * i + 1] * v[1] + M[3 * i + 2] * v[2]; }
void MatTVec3(const double* M, const double* v, double* r) { for (int i = 0; i < 3; ++i) r[i] = M[i] * v[0] + M[3 + i] * v[1] + M[6 + i] * v[2]; }

// pinhole K of beam_slam_launch/calibrations/ig2/cam.json (rectified), image 819 x 614
constexpr double kFx = 259.1959764852, kFy = 258.9512472321, kCx = 408.5323209475, kCy = 316.2596241094;
constexpr int kWidth = 819, kHeight = 614;
// T_imu_cam of bs_models/tests/reprojection_test.cpp:21-25; T_cam_baselink is its inverse
const double kT_imu_cam[16] = {0.0148655429818, -0.999880929698, 0.00414029679422, -0.0216401454975,
                               0.999557249008, 0.0149672133247, 0.025715529948, -0.064676986768,
                               -0.0257744366974, 0.00375618835797, 0.999660727178, 0.00981073058949, 0, 0, 0, 1};

void PerturbQuat(Rng& rng, double sigma_rad, const double q[4], double out[4]) {
  const double d[3] = {sigma_rad * rng.normal(), sigma_rad * rng.normal(), sigma_rad * rng.normal()};
  const double th = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  double dq[4] = {1, 0, 0, 0};
  if (th > 0) { const double k = std::sin(th / 2) / th; dq[0] = std::cos(th / 2); dq[1] = k * d[0]; dq[2] = k * d[1]; dq[3] = k * d[2]; }
  QuatMul(q, dq, out);
}

struct ConfigSpec { int kf, lm, track_lo, track_hi; bool vio, lidar, loops; int max_iters; double tol; };
ConfigSpec Spec(int config) {
  switch (config) {
    case 1: return {10, 300, 10, 10, true, false, false, 10, 1.5e-7};
    case 2: return {20, 1000, 14, 14, true, true, false, 10, 1.5e-7};
    case 3: return {200, 20000, 12, 17, false, false, false, 50, -1};
    case 4: return {2000, 200000, 12, 12, false, false, true, 50, -1};
    case 5: return {10000, 1000000, 12, 12, false, false, false, 50, -1};
    default: throw std::invalid_argument("config must be 1..5");
  }
}

}  // namespace

SolverOptions SyntheticSolverOptions(int config) {
  const ConfigSpec s = Spec(config);
  SolverOptions o;  // Ceres defaults
  o.max_num_iterations = s.max_iters;  // vio.yaml:13 / Ceres default 50
  if (s.tol > 0) { o.function_tolerance = o.gradient_tolerance = o.parameter_tolerance = s.tol; }  // vio.yaml:15-17
  return o;
}

SyntheticInfo BuildSynthetic(int config, double scale, uint64_t seed, HashGraph* graph, std::vector<std::pair<UUID, std::vector<double>>>* truth) {
  return BuildSyntheticTrajectory(config, scale, seed, 0, graph, truth);
}
double SyntheticKeyframeTime(int k) { return 1.0 + 0.7 * k; }
SyntheticInfo BuildSyntheticTrajectory(int config, double scale, uint64_t seed, int keyframes, HashGraph* graph, std::vector<std::pair<UUID, std::vector<double>>>* truth) {
  ConfigSpec spec = Spec(config);
  if (keyframes > 0) { spec.lm = static_cast<int>(std::lround(spec.lm * (double)keyframes / spec.kf)); spec.kf = keyframes; }
  SyntheticInfo info{};
  Rng rng(seed ? seed : 1000 + config);
  SeedConstraintUuids((seed ? seed : 1000 + config) * 7919 + 13);
  const int nkf = std::max(spec.vio ? 3 : 4, static_cast<int>(std::lround(spec.kf * (spec.vio ? 1.0 : scale))));
  const int nlm = std::max(8, static_cast<int>(std::lround(spec.lm * scale)));
  info.num_keyframes = nkf;
  graph->clear();

  double T_cam_baselink[16];
  Inverse(kT_imu_cam, 4, T_cam_baselink);
  const double K[9] = {kFx, 0, kCx, 0, kFy, kCy, 0, 0, 1};
  double Rcb[9], tcb[3];
  for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) Rcb[3 * i + j] = T_cam_baselink[4 * i + j]; tcb[i] = T_cam_baselink[4 * i + 3]; }

  // ---- keyframes ----
  struct Kf { double t; double R[9], q[4], p[3], v[3]; Variable::SharedPtr vq, vp, vv, vbg, vba; };
  std::vector<Kf> kfs(nkf);
  const double bg_true[3] = {rng.uniform(-1, 1) / 100, rng.uniform(-1, 1) / 100, rng.uniform(-1, 1) / 100};
  const double ba_true[3] = {rng.uniform(-1, 1) / 10, rng.uniform(-1, 1) / 10, rng.uniform(-1, 1) / 10};
  auto remember = [&](const Variable::SharedPtr& v, const double* tv) {
    if (truth) truth->emplace_back(v->uuid(), std::vector<double>(tv, tv + v->size()));
  };
  for (int k = 0; k < nkf; ++k) {
    Kf& f = kfs[k];
    f.t = SyntheticKeyframeTime(k);
    Attitude(f.t, f.R); RotToQuat(f.R, f.q); Position(f.t, f.p); Velocity(f.t, f.v);
    const Time stamp = Time::fromSec(f.t);
    f.vq = variables::Orientation3DStamped(stamp); f.vp = variables::Position3DStamped(stamp);
    remember(f.vq, f.q); remember(f.vp, f.p);
    // the first keyframe is the gauge anchor (window-start prior / fixed first pose): it starts at the truth
    const double sig = k == 0 ? 0.0 : 1.0;
    PerturbQuat(rng, sig * 0.017453292519943295, f.q, f.vq->data());                       // 1 deg
    for (int i = 0; i < 3; ++i) f.vp->data()[i] = f.p[i] + sig * 0.05 * rng.normal();     // 5 cm
    graph->addVariable(f.vq); graph->addVariable(f.vp);
    if (spec.vio) {
      f.vv = variables::VelocityLinear3DStamped(stamp); f.vbg = variables::GyroscopeBias3DStamped(stamp);
      f.vba = variables::AccelerationBias3DStamped(stamp);
      remember(f.vv, f.v); remember(f.vbg, bg_true); remember(f.vba, ba_true);
      for (int i = 0; i < 3; ++i) f.vv->data()[i] = f.v[i] + sig * 0.05 * rng.normal();   // 0.05 m/s; biases start at 0
      graph->addVariable(f.vv); graph->addVariable(f.vbg); graph->addVariable(f.vba);
    }
  }

  // ---- landmarks + reprojection constraints (visual_map.cpp:173-219: Cauchy(5*w), integer pixels) ----
  const double w_reproj = 1.0;
  int64_t nobs = 0; int made = 0;
  for (int l = 0; l < nlm; ++l) {
    for (int attempt = 0; attempt < 20; ++attempt) {
      const int len = std::min(nkf, rng.randint(spec.track_lo, spec.track_hi));
      const int s = rng.randint(0, nkf - len), m = s + len / 2;
      const double u = rng.uniform(20, kWidth - 20), v = rng.uniform(20, kHeight - 20), depth = rng.uniform(2, 30);
      double ray[3] = {(u - kCx) / kFx, (v - kCy) / kFy, 1.0};
      const double rn = std::sqrt(ray[0] * ray[0] + ray[1] * ray[1] + 1.0);
      double Pc[3] = {depth * ray[0] / rn, depth * ray[1] / rn, depth / rn};
      // P_b = Rcb^T (Pc - tcb); P_w = R_m P_b + p_m
      double d[3] = {Pc[0] - tcb[0], Pc[1] - tcb[1], Pc[2] - tcb[2]}, Pb[3], Pw[3];
      MatTVec3(Rcb, d, Pb); MatVec3(kfs[m].R, Pb, Pw);
      for (int i = 0; i < 3; ++i) Pw[i] += kfs[m].p[i];
      struct Ob { int k; double px[2]; };
      std::vector<Ob> obs;
      for (int k = s; k < s + len; ++k) {
        double dw[3] = {Pw[0] - kfs[k].p[0], Pw[1] - kfs[k].p[1], Pw[2] - kfs[k].p[2]}, pb[3], pc[3];
        MatTVec3(kfs[k].R, dw, pb); MatVec3(Rcb, pb, pc);
        for (int i = 0; i < 3; ++i) pc[i] += tcb[i];
        if (pc[2] < 0.5) continue;
        const double uu = kFx * pc[0] / pc[2] + kCx, vv = kFy * pc[1] / pc[2] + kCy;
        if (uu < 0 || uu >= kWidth || vv < 0 || vv >= kHeight) continue;
        obs.push_back({k, {std::round(uu + rng.normal()), std::round(vv + rng.normal())}});
      }
      if (obs.size() < 2) continue;
      auto lm = variables::Point3DLandmark(static_cast<uint64_t>(l));
      remember(lm, Pw);
      for (int i = 0; i < 3; ++i) lm->data()[i] = Pw[i] + 0.1 * rng.normal();  // 10 cm
      graph->addVariable(lm);
      for (const Ob& o : obs) {
        auto c = constraints::EuclideanReprojection("VO", *kfs[o.k].vq, *kfs[o.k].vp, *lm, T_cam_baselink, K, o.px, w_reproj);
        c->loss(CauchyLoss(5.0 * w_reproj));
        graph->addConstraint(c);
      }
      nobs += static_cast<int64_t>(obs.size()); ++made;
      break;
    }
  }
  info.num_landmarks = made; info.num_observations = nobs;

  if (spec.vio) {
    // ---- IMU preintegration between consecutive keyframes (imu_preintegration.cpp:246-320) ----
    const double h = 1.0 / 200.0, G[3] = {0, 0, -9.80665};
    const double zero3[3] = {0, 0, 0};
    for (int k = 0; k + 1 < nkf; ++k) {
      PreIntegrator pre; pre.Reset();
      const int nsamp = static_cast<int>(std::lround((kfs[k + 1].t - kfs[k].t) / h));
      for (int n = 0; n < nsamp; ++n) {
        IMUData d; d.t = kfs[k].t + n * h;
        const double tm = d.t + 0.5 * h;  // mid-interval sampling (imu_preintegration_tests.cpp:93-104)
        double R[9], a[3], w[3], f[3];
        Attitude(tm, R); Acceleration(tm, a); BodyRate(tm, w);
        const double aw[3] = {a[0] - G[0], a[1] - G[1], a[2] - G[2]};
        MatTVec3(R, aw, f);
        for (int i = 0; i < 3; ++i) { d.w[i] = w[i] + bg_true[i]; d.a[i] = f[i] + ba_true[i]; }
        pre.data.push_back(d);
      }
      pre.Integrate(kfs[k + 1].t, zero3, zero3, true, true, true);  // bias estimate at creation = 0
      double consts[288]; pre.FillFactorConstants(zero3, zero3, /*inertial_information_weight*/ 1.0, consts);
      const Variable* si[5] = {kfs[k].vq.get(), kfs[k].vp.get(), kfs[k].vv.get(), kfs[k].vbg.get(), kfs[k].vba.get()};
      const Variable* sj[5] = {kfs[k + 1].vq.get(), kfs[k + 1].vp.get(), kfs[k + 1].vv.get(), kfs[k + 1].vbg.get(), kfs[k + 1].vba.get()};
      graph->addConstraint(constraints::RelativeImuState3DStamped("IO", si, sj, consts));
      ++info.num_imu;
    }
    // ---- pseudo-marginalisation prior at the window start: mean = current state, cov = 1e-5 I
    //      (fixed_lag_smoother.cpp:261-267) ----
    double mean[16], cov[225] = {0};
    std::memcpy(mean, kfs[0].vq->data(), 32); std::memcpy(mean + 4, kfs[0].vp->data(), 24);
    std::memcpy(mean + 7, kfs[0].vv->data(), 24); std::memcpy(mean + 10, kfs[0].vbg->data(), 24);
    std::memcpy(mean + 13, kfs[0].vba->data(), 24);
    for (int i = 0; i < 15; ++i) cov[i * 15 + i] = 1e-5;
    const Variable* s0[5] = {kfs[0].vq.get(), kfs[0].vp.get(), kfs[0].vv.get(), kfs[0].vbg.get(), kfs[0].vba.get()};
    graph->addConstraint(constraints::AbsoluteImuState3DStamped("MARGINALIZATION", s0, mean, cov));
    ++info.num_priors;
  }

  auto relative_measurement = [&](const double* Ra, const double* pa, const double* Rb, const double* pb,
                                  double sig_p, double sig_r, double delta[7]) {
    double d[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]}, dp[3], RaT[9], Rab[9], q[4], qn[4];
    MatTVec3(Ra, d, dp); Transpose(Ra, 3, 3, RaT); MatMul(RaT, Rb, 3, 3, 3, Rab); RotToQuat(Rab, q);
    PerturbQuat(rng, sig_r, q, qn);
    for (int i = 0; i < 3; ++i) delta[i] = dp[i] + sig_p * rng.normal();
    for (int i = 0; i < 4; ++i) delta[3 + i] = qn[i];
  };
  double cov6[36] = {0};
  for (int i = 0; i < 6; ++i) cov6[i * 6 + i] = 1e-4;

  if (spec.lidar) {
    // ---- lidar odometry: relative pose between sensor frames with extrinsics
    //      (pose_3d_stamped_transaction.cpp:28-82; position free, orientation constant) ----
    auto pe = variables::Position3D("lidar_link", "base_link"); auto qe = variables::Orientation3D("lidar_link", "base_link");
    const double pe_true[3] = {0.10, 0.0, 0.20};
    double Re[9]; const double we[3] = {0.02, -0.01, 0.03}; ExpSO3(we, Re);
    double qe_true[4]; RotToQuat(Re, qe_true);
    remember(pe, pe_true); remember(qe, qe_true);
    for (int i = 0; i < 3; ++i) pe->data()[i] = pe_true[i] + 0.01 * rng.normal();
    for (int i = 0; i < 4; ++i) qe->data()[i] = qe_true[i];
    graph->addVariable(pe); graph->addVariable(qe);
    for (int k = 0; k + 1 < nkf; ++k) {
      double Ra[9], Rb[9], pa[3], pb[3], t[3];
      MatMul(kfs[k].R, Re, 3, 3, 3, Ra); MatVec3(kfs[k].R, pe_true, t); for (int i = 0; i < 3; ++i) pa[i] = t[i] + kfs[k].p[i];
      MatMul(kfs[k + 1].R, Re, 3, 3, 3, Rb); MatVec3(kfs[k + 1].R, pe_true, t); for (int i = 0; i < 3; ++i) pb[i] = t[i] + kfs[k + 1].p[i];
      double delta[7]; relative_measurement(Ra, pa, Rb, pb, 0.01, 0.002, delta);
      auto c = constraints::RelativePose3DStampedWithExtrinsics("LO", *kfs[k].vp, *kfs[k].vq, *kfs[k + 1].vp, *kfs[k + 1].vq, *pe, *qe, delta, cov6);
      c->loss(CauchyLoss(1.0));  // pose_3d_stamped_transaction.cpp:18
      graph->addConstraint(c); ++info.num_relative_pose;
    }
  }
  if (!spec.vio) {
    // ---- refinement graphs: odometry chain, optional loop closures, gauge prior on the first pose
    //      (submap_pose_graph_optimization.cpp:50-88,134-142) ----
    auto add_rel = [&](int a, int b) {
      double delta[7]; relative_measurement(kfs[a].R, kfs[a].p, kfs[b].R, kfs[b].p, 0.01, 0.002, delta);
      auto c = constraints::RelativePose3DStamped("PGO", *kfs[a].vp, *kfs[a].vq, *kfs[b].vp, *kfs[b].vq, delta, cov6);
      c->loss(CauchyLoss(1.0));
      graph->addConstraint(c); ++info.num_relative_pose;
    };
    for (int k = 0; k + 1 < nkf; ++k) add_rel(k, k + 1);
    if (spec.loops)
      for (int k = 120; k < nkf; k += 20) add_rel(k - 100 - rng.randint(0, 20), k);
    double mean[7], cov[36] = {0};
    std::memcpy(mean, kfs[0].vp->data(), 24); std::memcpy(mean + 3, kfs[0].vq->data(), 32);
    for (int i = 0; i < 6; ++i) cov[i * 6 + i] = 1e-9;  // pose_prior_noise_fixed_ (submap_pose_graph_optimization.h:40)
    graph->addConstraint(constraints::AbsolutePose3DStamped("PRIOR", *kfs[0].vp, *kfs[0].vq, mean, cov));
    ++info.num_priors;
  }
  info.num_variables = static_cast<int64_t>(graph->getVariables().size());
  info.num_constraints = static_cast<int64_t>(graph->getConstraints().size());
  return info;
}

}  // namespace bsfuse
