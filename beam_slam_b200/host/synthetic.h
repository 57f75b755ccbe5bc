#pragma once
#include <utility>
#include <vector>

#include "fuse_mirror.h"

namespace bsfuse {

struct SyntheticInfo {
  int num_keyframes, num_landmarks; long long num_observations; int num_imu, num_relative_pose, num_priors;
  long long num_variables, num_constraints;
};
// config = 1..5 (BASELINE.json configs in order); scale in (0,1] shrinks landmark (and, for 3-5, keyframe) counts
// for tests; seed 0 = 1000 + config. truth (optional) receives ground-truth values per variable UUID.
SyntheticInfo BuildSynthetic(int config, double scale, uint64_t seed, HashGraph* graph,
                             std::vector<std::pair<UUID, std::vector<double>>>* truth = nullptr);
// Same graph family over a longer trajectory: `keyframes` (> 0) replaces the config's keyframe count and scales the
// landmark count with it (source of the sliding-window replay, smoother_replay.h). Keyframe k is stamped
// SyntheticKeyframeTime(k).
SyntheticInfo BuildSyntheticTrajectory(int config, double scale, uint64_t seed, int keyframes, HashGraph* graph,
                                       std::vector<std::pair<UUID, std::vector<double>>>* truth = nullptr);
double SyntheticKeyframeTime(int k);
// Solver options the survey names per config (vio.yaml:13-17 for 1-2, Ceres defaults for 3-5).
SolverOptions SyntheticSolverOptions(int config);

}  // namespace bsfuse
