#include "smoother_replay.h"

#include <chrono>
#include <cstring>

#include "synthetic.h"

namespace bsfuse {

namespace {
const char* kStateTypes[5] = {"fuse_variables::Orientation3DStamped", "fuse_variables::Position3DStamped",
                              "fuse_variables::VelocityLinear3DStamped", "bs_variables::GyroscopeBias3DStamped",
                              "bs_variables::AccelerationBias3DStamped"};
}

SmootherReplay::SmootherReplay(int total_keyframes, int window_keyframes, uint64_t seed)
    : total_(total_keyframes), window_kf_(window_keyframes), next_(0) {
  if (window_keyframes < 2 || total_keyframes < window_keyframes) throw std::invalid_argument("SmootherReplay: need total >= window >= 2 keyframes");
  BuildSyntheticTrajectory(1, 1.0, seed, total_keyframes, &source_);
  kf_vars_.resize(total_); kf_constraints_.resize(total_);
  std::unordered_map<UUID, int, UuidHash> kf_of;
  for (int k = 0; k < total_; ++k) {
    const Time stamp = Time::fromSec(SyntheticKeyframeTime(k));
    for (const char* type : kStateTypes) {
      const UUID u = uuid::generate(type, stamp.sec, stamp.nsec, uuid::nil());
      kf_vars_[k].push_back(u); kf_of[u] = k;
    }
  }
  for (const auto& kv : source_.getConstraints()) {
    if (kv.second->source() == "MARGINALIZATION") continue;   // the window start prior is rebuilt every cycle
    int newest = -1;
    for (const UUID& u : kv.second->variables()) { auto it = kf_of.find(u); if (it != kf_of.end()) newest = std::max(newest, it->second); }
    if (newest >= 0) kf_constraints_[newest].push_back(kv.second);
  }
  for (int k = 0; k < window_kf_; ++k) AddKeyframe(k);
  next_ = window_kf_;
  MarginaliseKeyframe(-1);   // prior on the first window state
}

void SmootherReplay::AddKeyframe(int k) {
  for (const UUID& u : kf_vars_[k]) window_.addVariable(source_.getVariable(u).clone());
  for (const auto& c : kf_constraints_[k]) {
    for (const UUID& u : c->variables()) if (!window_.variableExists(u)) window_.addVariable(source_.getVariable(u).clone());   // landmarks
    window_.addConstraint(std::make_shared<Constraint>(*c));   // fuse clones constraints into the graph (graph.cpp:69-95)
  }
}

void SmootherReplay::MarginaliseKeyframe(int k) {
  if (k >= 0) {
    std::vector<UUID> touched;
    for (const UUID& u : kf_vars_[k]) {
      for (const UUID& c : window_.getConnectedConstraints(u)) {
        if (!window_.constraintExists(c)) continue;
        for (const UUID& v : window_.getConstraint(c).variables()) touched.push_back(v);
        window_.removeConstraint(c);
      }
    }
    for (const UUID& u : kf_vars_[k]) window_.removeVariable(u);
    for (const UUID& v : touched)   // landmarks nobody observes any more leave with the keyframe
      if (window_.variableExists(v) && window_.getConnectedConstraints(v).empty()) window_.removeVariable(v);
  }
  const int first = k + 1;
  const Variable* s0[5]; double mean[16], cov[225] = {0};
  int o = 0;
  for (int i = 0; i < 5; ++i) { s0[i] = &window_.getVariable(kf_vars_[first][i]); std::memcpy(mean + o, s0[i]->data(), sizeof(double) * s0[i]->size()); o += static_cast<int>(s0[i]->size()); }
  for (int i = 0; i < 15; ++i) cov[i * 15 + i] = 1e-5;
  window_.addConstraint(constraints::AbsoluteImuState3DStamped("MARGINALIZATION", s0, mean, cov));
}

double SmootherReplay::Advance() {
  if (Done()) throw std::logic_error("SmootherReplay: trajectory exhausted");
  const auto t0 = std::chrono::steady_clock::now();
  AddKeyframe(next_);
  MarginaliseKeyframe(next_ - window_kf_);
  ++next_;
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
SolverSummary SmootherReplay::Step(const SolverOptions& options, double* update_seconds) {
  const double t = Advance();
  if (update_seconds) *update_seconds = t;
  return window_.optimize(options);
}

long long SmootherReplay::num_observations() const {
  long long n = 0;
  for (const auto& kv : window_.getConstraints()) n += kv.second->type() == "bs_constraints::EuclideanReprojectionConstraint";
  return n;
}

}  // namespace bsfuse
