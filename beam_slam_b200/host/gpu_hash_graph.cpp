#include "gpu_hash_graph.h"

#include <chrono>

namespace bsfuse {

void ToCOptions(const SolverOptions& o, bsnls_options* c) {
  bsnls_default_options(c);
  c->max_num_iterations = o.max_num_iterations;
  c->max_solver_time_in_seconds = o.max_solver_time_in_seconds;
  c->function_tolerance = o.function_tolerance; c->gradient_tolerance = o.gradient_tolerance;
  c->parameter_tolerance = o.parameter_tolerance;
  c->initial_trust_region_radius = o.initial_trust_region_radius;
  c->max_trust_region_radius = o.max_trust_region_radius; c->min_trust_region_radius = o.min_trust_region_radius;
  c->min_relative_decrease = o.min_relative_decrease;
  c->min_lm_diagonal = o.min_lm_diagonal; c->max_lm_diagonal = o.max_lm_diagonal;
  c->max_num_consecutive_invalid_steps = o.max_num_consecutive_invalid_steps;
  c->jacobi_scaling = o.jacobi_scaling ? 1 : 0;
  c->linear_solver = o.linear_solver;
}

GpuHashGraph::GpuHashGraph() = default;
GpuHashGraph::~GpuHashGraph() { if (handle_) bsnls_destroy(handle_); }

void GpuHashGraph::EnsureHandle(const SolverOptions& options) const {
  if (handle_ && options.linear_solver != last_options_.linear_solver) device_version_ = 0;   // reduced-system storage changes
  last_options_ = options;
  bsnls_options c; ToCOptions(options, &c);
  if (!handle_) {
    if (bsnls_create(&c, &handle_) != BSNLS_OK) {
      const char* e = bsnls_last_error(nullptr);
      throw std::runtime_error(std::string("bsnls_create failed: ") + (e ? e : "?"));
    }
  } else if (bsnls_set_options(handle_, &c) != BSNLS_OK) {
    throw std::invalid_argument(bsnls_last_error(handle_));
  }
}

const FlatProblem& GpuHashGraph::flattened() const {
  if (flat_version_ == structureVersion() && flat_version_ != 0) flat_.GatherValues();   // only Variable::data() moved
  else { Flatten(*this, &flat_); flat_version_ = structureVersion(); }                 // createProblem
  return flat_;
}

int GpuHashGraph::SyncProblem() const {
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  flattened();
  const auto t1 = clk::now();
  timing_.flatten = std::chrono::duration<double>(t1 - t0).count();
  int rc = BSNLS_OK;
  if (flat_.cgroups.empty() || flat_.block_size.empty()) { device_version_ = 0; timing_.upload = 0; timing_.structure_reused = false; return BSNLS_OK; }
  if (device_version_ == flat_version_ && device_version_ != 0) {
    // same blocks, same factors (constraints are immutable once added, as in fuse): the device problem stays, values move
    rc = bsnls_set_values(handle_, flat_.block_values.data());
    timing_.structure_reused = true;
  } else {
    rc = bsnls_set_problem(handle_, &flat_.c);
    device_version_ = rc == BSNLS_OK ? flat_version_ : 0;
    timing_.structure_reused = false;
  }
  timing_.upload = std::chrono::duration<double>(clk::now() - t1).count();
  return rc;
}

SolverSummary GpuHashGraph::optimize(const SolverOptions& options) {
  const auto t0 = std::chrono::steady_clock::now();
  std::string err;
  if (!options.IsValid(&err)) throw std::invalid_argument("Invalid solver options: " + err);
  SolverSummary s;
  if (getConstraints().empty() || getVariables().empty()) {  // empty problem: Ceres converges at once
    s.termination_type = SolverSummary::CONVERGENCE; s.message = "Function tolerance reached. No non-constant parameter blocks found.";
    return s;
  }
  EnsureHandle(options);
  bsnls_summary cs{};
  std::vector<bsnls_iteration> its(static_cast<size_t>(options.max_num_iterations) + 2);
  int rc = SyncProblem();
  const auto t1 = std::chrono::steady_clock::now();
  if (rc == BSNLS_OK && options.max_solver_time_in_seconds < 1e8) {
    // optimizeFor: what createProblem's counterpart took is charged against the budget (hash_graph.cpp:433-439)
    bsnls_options c; ToCOptions(options, &c);
    c.max_solver_time_in_seconds = std::max(0.0, options.max_solver_time_in_seconds - std::chrono::duration<double>(t1 - t0).count());
    rc = bsnls_set_options(handle_, &c);
  }
  if (rc == BSNLS_OK) rc = bsnls_solve(handle_, &cs, its.data(), static_cast<int32_t>(its.size()));
  const auto t2 = std::chrono::steady_clock::now();
  timing_.solve = std::chrono::duration<double>(t2 - t1).count(); timing_.download = 0;
  if (rc != BSNLS_OK) {
    s.termination_type = SolverSummary::FAILURE;
    s.message = std::string("bsnls: ") + bsnls_last_error(handle_);
    return s;  // variables untouched; caller sees !IsSolutionUsable() (fixed_lag_smoother.cpp:286)
  }
  s.termination_type = static_cast<SolverSummary::TerminationType>(cs.termination_type);
  s.message = cs.message;
  s.initial_cost = cs.initial_cost; s.final_cost = cs.final_cost; s.fixed_cost = cs.fixed_cost;
  s.device_time_in_seconds = cs.device_time_in_seconds;
  s.num_successful_steps = cs.num_successful_steps; s.num_unsuccessful_steps = cs.num_unsuccessful_steps;
  s.num_e_blocks = cs.num_e_blocks; s.num_f_blocks = cs.num_f_blocks; s.reduced_system_dim = cs.reduced_system_dim;
  s.linear_solver_used = cs.linear_solver_used; s.reduced_system_nnz_blocks = cs.reduced_system_nnz_blocks;
  for (int i = 0; i < cs.num_iterations && i < static_cast<int>(its.size()); ++i) {
    IterationSummary it;
    it.iteration = its[i].iteration; it.step_is_valid = its[i].step_is_valid; it.step_is_successful = its[i].step_is_successful;
    it.linear_solver_iterations = its[i].linear_solver_iterations;
    it.cost = its[i].cost; it.cost_change = its[i].cost_change; it.gradient_max_norm = its[i].gradient_max_norm;
    it.gradient_norm = its[i].gradient_norm; it.step_norm = its[i].step_norm;
    it.relative_decrease = its[i].relative_decrease; it.trust_region_radius = its[i].trust_region_radius;
    s.iterations.push_back(it);
  }
  if (s.IsSolutionUsable()) {
    if (bsnls_get_values(handle_, flat_.block_values.data()) != BSNLS_OK) {
      s.termination_type = SolverSummary::FAILURE; s.message = std::string("bsnls: ") + bsnls_last_error(handle_);
      return s;
    }
    flat_.ScatterValues(flat_.block_values.data());  // Ceres' in-place update of Variable::data()
    timing_.download = std::chrono::duration<double>(std::chrono::steady_clock::now() - t2).count();
  }
  s.total_time_in_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return s;
}

void GpuHashGraph::Linearize(const std::vector<const Constraint*>& constraints, std::vector<double>* residuals, std::vector<double>* jacobians) const {
  EnsureHandle(last_options_);
  if (SyncProblem() != BSNLS_OK) throw std::runtime_error(std::string("linearize: ") + bsnls_last_error(handle_));
  std::vector<int32_t> ids;
  for (const Constraint* c : constraints) {
    if (c->flat_hint.owner != &flat_ || c->flat_hint.res_index < 0) throw std::logic_error("linearize: the constraint is not part of this graph");
    ids.push_back(c->flat_hint.res_index);
  }
  residuals->assign(ids.size() * BSNLS_MARGINAL_MAX_ROWS, 0.0); jacobians->assign(ids.size() * BSNLS_MARGINAL_MAX_ROWS * 30, 0.0);
  if (bsnls_linearize(handle_, static_cast<int32_t>(ids.size()), ids.data(), residuals->data(), jacobians->data()) != BSNLS_OK)
    throw std::runtime_error(std::string("linearize: ") + bsnls_last_error(handle_));
}

SolverSummary GpuHashGraph::optimizeFor(double max_time_s, const SolverOptions& options) {
  // The reference subtracts the createProblem time from the budget (hash_graph.cpp:433-439): optimize() does the same
  // with the time flattening and bsnls_set_problem took before it hands the rest to the device loop.
  SolverOptions o = options;
  o.max_solver_time_in_seconds = std::max(0.0, max_time_s);
  return optimize(o);
}

void GpuHashGraph::getCovariance(const std::vector<std::pair<UUID, UUID>>& requests, std::vector<std::vector<double>>& matrices,
                                 bool use_tangent_space) const {
  if (requests.empty()) return;   // "Avoid doing a bunch of work if the request is empty" (hash_graph.cpp:309-313)
  for (const auto& r : requests) { getVariable(r.first); getVariable(r.second); }   // std::out_of_range, as :337-349
  EnsureHandle(last_options_);
  if (SyncProblem() != BSNLS_OK) throw std::runtime_error(std::string("Could not compute requested covariance blocks: ") + bsnls_last_error(handle_));
  std::vector<int32_t> pairs; pairs.reserve(2 * requests.size());
  for (const auto& r : requests) { pairs.push_back(getVariable(r.first).flat_index); pairs.push_back(getVariable(r.second).flat_index); }
  std::vector<double> out(16 * requests.size());
  if (bsnls_get_covariance(handle_, pairs.data(), static_cast<int32_t>(requests.size()), use_tangent_space ? 1 : 0, out.data()) != BSNLS_OK)
    throw std::runtime_error(std::string("Could not compute requested covariance blocks: ") + bsnls_last_error(handle_));
  matrices.resize(requests.size());
  for (size_t i = 0; i < requests.size(); ++i) {
    const Variable& a = getVariable(requests[i].first); const Variable& b = getVariable(requests[i].second);
    const size_t n = use_tangent_space ? a.localSize() * b.localSize() : a.size() * b.size();
    matrices[i].assign(out.begin() + 16 * i, out.begin() + 16 * i + n);
  }
}

bool GpuHashGraph::evaluate(double* cost, std::vector<double>* residuals, std::vector<double>* gradient) const {
  if (getConstraints().empty() || getVariables().empty()) { if (cost) *cost = 0.0; if (residuals) residuals->clear(); if (gradient) gradient->clear(); return true; }
  EnsureHandle(last_options_);
  if (SyncProblem() != BSNLS_OK) return false;
  int64_t nr = 0, ng = 0;
  if (bsnls_evaluate(handle_, cost, nullptr, &nr, nullptr, &ng) != BSNLS_OK) return false;
  if (residuals) residuals->assign(nr, 0.0);
  if (gradient) gradient->assign(ng, 0.0);
  if (residuals || gradient)
    return bsnls_evaluate(handle_, cost, residuals ? residuals->data() : nullptr, &nr,
                          gradient ? gradient->data() : nullptr, &ng) == BSNLS_OK;
  return true;
}

}  // namespace bsfuse
