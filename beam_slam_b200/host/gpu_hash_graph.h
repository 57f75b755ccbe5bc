// GpuHashGraph — the drop-in: a HashGraph whose optimize / optimizeFor / evaluate run on the GPU through the
// C-ABI of include/bsnls.h instead of ceres::Solve (fuse_graphs/src/hash_graph.cpp:416-455). In a fuse workspace
// the same class derives from fuse_graphs::HashGraph and is exported with
// PLUGINLIB_EXPORT_CLASS(..., fuse_core::Graph); containers and createProblem order are inherited.
#pragma once
#include "flatten.h"

namespace bsfuse {

class GpuHashGraph : public HashGraph {
 public:
  GpuHashGraph();
  ~GpuHashGraph() override;
  GpuHashGraph(const GpuHashGraph&) = delete;
  GpuHashGraph& operator=(const GpuHashGraph&) = delete;
  // ceres::Solver::Summary optimize(const ceres::Solver::Options&) — variables are updated in place unless the
  // solve FAILS; invalid options throw std::invalid_argument (ceres_options.cpp:195-199).
  SolverSummary optimize(const SolverOptions& options = SolverOptions()) override;
  // optimizeFor(max_optimization_time, options): flattening time is charged against the budget (hash_graph.cpp:428-444)
  SolverSummary optimizeFor(double max_optimization_time_s, const SolverOptions& options = SolverOptions()) override;
  // Problem::Evaluate with default EvaluateOptions (hash_graph.cpp:446-455)
  bool evaluate(double* cost, std::vector<double>* residuals, std::vector<double>* gradient) const override;
  const FlatProblem& lastProblem() const { return flat_; }
 private:
  void EnsureHandle(const SolverOptions& options) const;
  mutable bsnls_handle* handle_ = nullptr;
  mutable FlatProblem flat_;
};

void ToCOptions(const SolverOptions& o, bsnls_options* c);

}  // namespace bsfuse
