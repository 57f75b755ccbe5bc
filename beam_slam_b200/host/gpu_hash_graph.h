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
  // ceres::Covariance on the GPU: blocks of the inverse normal matrix at the current values (hash_graph.cpp:303-414).
  // Landmark blocks are not supported (std::runtime_error), as are constant blocks (Ceres has no covariance for them).
  void getCovariance(const std::vector<std::pair<UUID, UUID>>& covariance_requests,
                     std::vector<std::vector<double>>& covariance_matrices, bool use_tangent_space = true) const override;
  const FlatProblem& lastProblem() const { return flat_; }
  // createProblem-order SoA view of the graph as it is now (re-flattened only if the structure changed, else the values
  // are gathered); the returned object stays owned by the graph
  const FlatProblem& flattened() const;
  // host wall-clock split of the last optimize(), seconds: flatten (or value gather), bsnls_set_problem (or
  // bsnls_set_values), bsnls_solve, bsnls_get_values + scatter; and whether the device problem was reused
  struct Timing { double flatten = 0, upload = 0, solve = 0, download = 0; bool structure_reused = false; };
  const Timing& lastTiming() const { return timing_; }
  // fuse_constraints::detail::linearize (marginalize_variables.cpp:281-399) of constraints of this graph at the current
  // values, on the device (bsnls_linearize): per constraint BSNLS_MARGINAL_MAX_ROWS residuals and a
  // BSNLS_MARGINAL_MAX_ROWS x 30 row-major tangent Jacobian, three columns per variable of the constraint.
  // std::runtime_error if the device call fails.
  void Linearize(const std::vector<const Constraint*>& constraints, std::vector<double>* residuals, std::vector<double>* jacobians) const;
 private:
  void EnsureHandle(const SolverOptions& options) const;
  // Brings the device problem up to date with the graph: unchanged structure (structureVersion) -> values only;
  // otherwise re-flatten (cached per-constraint lookups) and bsnls_set_problem. Returns a bsnls status.
  int SyncProblem() const;
  mutable bsnls_handle* handle_ = nullptr;
  mutable FlatProblem flat_;
  mutable uint64_t flat_version_ = 0;     // structureVersion() flat_ was built from (0: none)
  mutable uint64_t device_version_ = 0;   // structureVersion() the device problem was built from (0: none)
  mutable Timing timing_;
  mutable SolverOptions last_options_;   // evaluate() keeps the handle's options (re-capturing CUDA graphs costs more than a solve)
};

void ToCOptions(const SolverOptions& o, bsnls_options* c);

}  // namespace bsfuse
