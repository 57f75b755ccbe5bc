// Replay of the fixed-lag smoother's cycle on a synthetic VIO trajectory: the call pattern the drop-in graph sees from
// bs_optimizers::FixedLagSmoother::optimizationLoop (bs_optimizers/src/fixed_lag_smoother.cpp:166-311). Every cycle
//   1. graph->update(new transaction): the next keyframe's five state variables, the landmarks it sees for the first
//      time, its reprojection constraints and the IMU-preintegration constraint from the previous keyframe (:220);
//   2. pseudo-marginalisation of the keyframe that left the window: its variables and every constraint connected to
//      them are removed, landmarks left without constraints go with them, and an AbsoluteImuState3DStampedConstraint
//      with covariance 1e-5 I is put on the new window start at its current estimate (:244-268, :742-797);
//   3. graph->optimize(solver_options) (:281).
// Used by bench.py and the tests to time and check optimize() when the graph's structure changes every call.
#pragma once
#include <vector>

#include "gpu_hash_graph.h"

namespace bsfuse {

class SmootherReplay {
 public:
  // trajectory of `total_keyframes` VIO keyframes (config-1 family: ~30 landmarks first seen per keyframe, tracks of 10),
  // window of `window_keyframes`
  SmootherReplay(int total_keyframes, int window_keyframes, uint64_t seed);
  bool Done() const { return next_ >= total_; }
  // steps 1 and 2 of a cycle; returns the host time they took, seconds
  double Advance();
  // one cycle; returns optimize()'s summary. update_seconds (optional) = host time of steps 1 and 2.
  SolverSummary Step(const SolverOptions& options, double* update_seconds = nullptr);
  GpuHashGraph& graph() { return window_; }
  int window_start() const { return next_ - window_kf_; }
  long long num_observations() const;
 private:
  void AddKeyframe(int k);
  void MarginaliseKeyframe(int k);
  GpuHashGraph source_, window_;   // source_ is only a container for the whole trajectory's variables and constraints
  int total_, window_kf_, next_;
  std::vector<std::vector<UUID>> kf_vars_;                       // per keyframe: q, p, v, bg, ba
  std::vector<std::vector<Constraint::SharedPtr>> kf_constraints_;   // constraints whose newest keyframe is k
};

}  // namespace bsfuse
