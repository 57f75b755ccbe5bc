#include "fuse_mirror.h"

#include <cmath>
#include <cstdio>
#include <mutex>
#include <sstream>

#include "linalg.h"

namespace bsfuse {

static uuid::RandomGenerator& Gen() { static uuid::RandomGenerator g(0x5eed); return g; }
static std::mutex g_gen_mutex;
UUID NextConstraintUuid() { std::lock_guard<std::mutex> l(g_gen_mutex); return Gen()(); }
void SeedConstraintUuids(uint64_t seed) { std::lock_guard<std::mutex> l(g_gen_mutex); Gen() = uuid::RandomGenerator(seed); }

// covariance.inverse().llt().matrixU(), row-major (fuse_core::MatrixNd are row-major, fuse_core/eigen.h:59-69)
bool SqrtInformationFromCovariance(const double* cov, int n, double* A) {
  std::vector<double> inv(n * n), L(n * n);
  if (!linalg::Inverse(cov, n, inv.data())) return false;
  if (!linalg::CholeskyLower(inv.data(), n, L.data())) return false;
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) A[i * n + j] = (j >= i) ? L[j * n + i] : 0.0;
  return true;
}

namespace constraints {

static void Require(bool ok, const char* what) { if (!ok) throw std::invalid_argument(what); }

Constraint::SharedPtr EuclideanReprojection(const std::string& source, const Variable& q, const Variable& t,
    const Variable& P, const double T[16], const double K[9], const double pixel[2], double w) {
  std::vector<double> pl = {pixel[0], pixel[1], K[0], K[4], K[2], K[5]};
  for (int i = 0; i < 12; ++i) pl.push_back(T[i]);
  pl.push_back(w);
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::EuclideanReprojectionConstraint", source,
                                      std::vector<UUID>{q.uuid(), t.uuid(), P.uuid()}, pl);
}
Constraint::SharedPtr RelativeImuState3DStamped(const std::string& source, const Variable* const si[5],
    const Variable* const sj[5], const double payload[288]) {
  std::vector<UUID> v;
  for (int k = 0; k < 5; ++k) v.push_back(si[k]->uuid());
  for (int k = 0; k < 5; ++k) v.push_back(sj[k]->uuid());
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::RelativeImuState3DStampedConstraint", source,
                                      v, std::vector<double>(payload, payload + 288));
}
Constraint::SharedPtr AbsoluteImuState3DStamped(const std::string& source, const Variable* const s[5],
    const double mean[16], const double cov[225]) {
  std::vector<double> pl(mean, mean + 16); pl.resize(16 + 225);
  Require(SqrtInformationFromCovariance(cov, 15, pl.data() + 16), "AbsoluteImuState3DStampedConstraint: covariance not SPD");
  std::vector<UUID> v; for (int k = 0; k < 5; ++k) v.push_back(s[k]->uuid());
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::AbsoluteImuState3DStampedConstraint", source, v, pl);
}
static std::vector<double> PosePayload(const double d[7], const double cov[36], const char* who) {
  std::vector<double> pl(d, d + 7); pl.resize(7 + 36);
  Require(SqrtInformationFromCovariance(cov, 6, pl.data() + 7), who);
  return pl;
}
Constraint::SharedPtr RelativePose3DStamped(const std::string& source, const Variable& p1, const Variable& q1,
    const Variable& p2, const Variable& q2, const double delta[7], const double cov[36]) {
  return std::make_shared<Constraint>(NextConstraintUuid(), "fuse_constraints::RelativePose3DStampedConstraint", source,
      std::vector<UUID>{p1.uuid(), q1.uuid(), p2.uuid(), q2.uuid()}, PosePayload(delta, cov, "RelativePose3DStampedConstraint: covariance not SPD"));
}
Constraint::SharedPtr RelativePose3DStampedWithExtrinsics(const std::string& source, const Variable& p1,
    const Variable& q1, const Variable& p2, const Variable& q2, const Variable& pe, const Variable& qe,
    const double delta[7], const double cov[36]) {
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::RelativePose3DStampedWithExtrinsicsConstraint", source,
      std::vector<UUID>{p1.uuid(), q1.uuid(), p2.uuid(), q2.uuid(), pe.uuid(), qe.uuid()},
      PosePayload(delta, cov, "RelativePose3DStampedWithExtrinsicsConstraint: covariance not SPD"));
}
Constraint::SharedPtr AbsolutePose3DStamped(const std::string& source, const Variable& p, const Variable& q,
    const double mean[7], const double cov[36]) {
  return std::make_shared<Constraint>(NextConstraintUuid(), "fuse_constraints::AbsolutePose3DStampedConstraint", source,
      std::vector<UUID>{p.uuid(), q.uuid()}, PosePayload(mean, cov, "AbsolutePose3DStampedConstraint: covariance not SPD"));
}
Constraint::SharedPtr RelativeOrientation3DStamped(const std::string& source, const Variable& q1, const Variable& q2,
    const double delta[4], const double cov[9]) {
  std::vector<double> pl(delta, delta + 4); pl.resize(13);
  Require(SqrtInformationFromCovariance(cov, 3, pl.data() + 4), "RelativeOrientation3DStampedConstraint: covariance not SPD");
  return std::make_shared<Constraint>(NextConstraintUuid(), "fuse_constraints::RelativeOrientation3DStampedConstraint", source,
      std::vector<UUID>{q1.uuid(), q2.uuid()}, pl);
}
Constraint::SharedPtr AbsoluteOrientation3DStamped(const std::string& source, const Variable& q, const double mean[4], const double cov[9]) {
  std::vector<double> pl(mean, mean + 4); pl.resize(13);
  Require(SqrtInformationFromCovariance(cov, 3, pl.data() + 4), "AbsoluteOrientation3DStampedConstraint: covariance not SPD");
  return std::make_shared<Constraint>(NextConstraintUuid(), "fuse_constraints::AbsoluteOrientation3DStampedConstraint", source,
      std::vector<UUID>{q.uuid()}, pl);
}
Constraint::SharedPtr GravityAlignmentStamped(const std::string& source, const Variable& q, const double gravity_in_baselink[3], const double cov[4]) {
  std::vector<double> pl(gravity_in_baselink, gravity_in_baselink + 3); pl.resize(7);
  Require(SqrtInformationFromCovariance(cov, 2, pl.data() + 3), "GravityAlignmentStampedConstraint: covariance not SPD");
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::GravityAlignmentStampedConstraint", source,
      std::vector<UUID>{q.uuid()}, pl);
}
static std::vector<double> Vec3Payload(const double m[3], const double cov[9]) {
  std::vector<double> pl(m, m + 3); pl.resize(12);
  Require(SqrtInformationFromCovariance(cov, 3, pl.data() + 3), "vector constraint: covariance not SPD");
  return pl;
}
Constraint::SharedPtr AbsoluteVec3(const std::string& source, const Variable& x, const double mean[3], const double cov[9]) {
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::AbsoluteConstraint3D", source,
      std::vector<UUID>{x.uuid()}, Vec3Payload(mean, cov));
}
Constraint::SharedPtr RelativeVec3(const std::string& source, const Variable& x1, const Variable& x2,
    const double delta[3], const double cov[9]) {
  return std::make_shared<Constraint>(NextConstraintUuid(), "bs_constraints::RelativeConstraint3D", source,
      std::vector<UUID>{x1.uuid(), x2.uuid()}, Vec3Payload(delta, cov));
}
}  // namespace constraints

bool SolverOptions::IsValid(std::string* error) const {
  auto bad = [&](const char* m) { if (error) *error = m; return false; };
  if (max_num_iterations < 0) return bad("max_num_iterations >= 0 violated");
  if (max_solver_time_in_seconds < 0) return bad("max_solver_time_in_seconds >= 0 violated");
  if (function_tolerance < 0) return bad("function_tolerance >= 0 violated");
  if (gradient_tolerance < 0) return bad("gradient_tolerance >= 0 violated");
  if (parameter_tolerance < 0) return bad("parameter_tolerance >= 0 violated");
  if (!(initial_trust_region_radius > 0)) return bad("initial_trust_region_radius > 0 violated");
  if (!(min_trust_region_radius > 0)) return bad("min_trust_region_radius > 0 violated");
  if (!(max_trust_region_radius > 0)) return bad("max_trust_region_radius > 0 violated");
  if (min_trust_region_radius > max_trust_region_radius) return bad("min_trust_region_radius <= max_trust_region_radius violated");
  if (min_trust_region_radius > initial_trust_region_radius) return bad("min_trust_region_radius <= initial_trust_region_radius violated");
  if (initial_trust_region_radius > max_trust_region_radius) return bad("initial_trust_region_radius <= max_trust_region_radius violated");
  if (min_relative_decrease < 0) return bad("min_relative_decrease >= 0 violated");
  if (min_lm_diagonal < 0 || max_lm_diagonal < 0 || min_lm_diagonal > max_lm_diagonal) return bad("lm_diagonal bounds violated");
  if (max_num_consecutive_invalid_steps < 0) return bad("max_num_consecutive_invalid_steps >= 0 violated");
  if (num_threads <= 0) return bad("num_threads > 0 violated");
  return true;
}

std::string SolverSummary::BriefReport() const {
  char buf[256];
  static const char* names[] = {"CONVERGENCE", "NO_CONVERGENCE", "FAILURE"};
  std::snprintf(buf, sizeof(buf), "bsnls Report: Iterations: %d, Initial cost: %e, Final cost: %e, Termination: %s",
                static_cast<int>(iterations.size()), initial_cost, final_cost, names[termination_type]);
  return buf;
}
std::string SolverSummary::FullReport() const {
  std::ostringstream o;
  o << BriefReport() << "\n" << message << "\n"
    << "e-blocks " << num_e_blocks << ", f-blocks " << num_f_blocks << ", reduced system " << reduced_system_dim
    << " (" << (linear_solver_used == 0 ? "dense Cholesky" : "block-sparse PCG") << ")\n"
    << "iter      cost      cost_change  |gradient|   |step|    tr_ratio  tr_radius\n";
  char buf[200];
  for (const auto& it : iterations) {
    std::snprintf(buf, sizeof(buf), "%4d  %.6e  %9.2e  %9.2e  %9.2e  %9.2e  %9.2e\n", it.iteration, it.cost,
                  it.cost_change, it.gradient_max_norm, it.step_norm, it.relative_decrease, it.trust_region_radius);
    o << buf;
  }
  o << "total " << total_time_in_seconds << " s, device " << device_time_in_seconds << " s\n";
  return o.str();
}

// ---- HashGraph containers: semantics of fuse_graphs/src/hash_graph.cpp:100-301 ----
void HashGraph::clear() { constraints_.clear(); variables_.clear(); variables_on_hold_.clear(); constraints_by_variable_uuid_.clear(); ++structure_version_; }

bool HashGraph::addConstraint(Constraint::SharedPtr c) {
  if (!c) return false;
  if (constraintExists(c->uuid())) return false;
  for (const auto& v : c->variables())
    if (!variableExists(v))
      throw std::logic_error("Attempting to add a constraint (" + uuid::to_string(c->uuid()) +
                             ") that uses an unknown variable (" + uuid::to_string(v) + ").");
  for (const auto& v : c->variables()) constraints_by_variable_uuid_[v].insert(c->uuid());
  constraints_.emplace(c->uuid(), std::move(c));
  ++structure_version_;
  return true;
}
bool HashGraph::removeConstraint(const UUID& u) {
  auto it = constraints_.find(u);
  if (it == constraints_.end()) return false;
  for (const auto& v : it->second->variables()) constraints_by_variable_uuid_[v].erase(u);
  constraints_.erase(it);
  ++structure_version_;
  return true;
}
const Constraint& HashGraph::getConstraint(const UUID& u) const {
  auto it = constraints_.find(u);
  if (it == constraints_.end()) throw std::out_of_range("The constraint UUID " + uuid::to_string(u) + " does not exist.");
  return *it->second;
}
std::vector<UUID> HashGraph::getConnectedConstraints(const UUID& variable) const {
  std::vector<UUID> out;
  auto it = constraints_by_variable_uuid_.find(variable);
  if (it != constraints_by_variable_uuid_.end()) out.assign(it->second.begin(), it->second.end());
  else if (!variableExists(variable)) throw std::logic_error("Attempting to access constraints connected to variable (" + uuid::to_string(variable) + "), but that variable does not exist in this graph.");
  return out;
}
bool HashGraph::addVariable(Variable::SharedPtr v) {
  if (!v) return false;
  auto it = variables_.find(v->uuid());
  if (it != variables_.end()) {
    // update in place (hash_graph.cpp:229-238). The stored object keeps its identity (constraints of this graph cache
    // pointers to it); only a change of shape or hold flag is a structural change.
    Variable& old = *it->second;
    if (old.size() != v->size() || old.manifold() != v->manifold() || old.holdConstant() != v->holdConstant() || old.type() != v->type()) ++structure_version_;
    const int32_t keep = old.flat_index;
    old = *v; old.flat_index = keep;
    return true;
  }
  variables_.emplace(v->uuid(), std::move(v));
  ++structure_version_;
  return true;
}
bool HashGraph::removeVariable(const UUID& u) {
  auto it = variables_.find(u);
  if (it == variables_.end()) return false;
  auto c = constraints_by_variable_uuid_.find(u);
  if (c != constraints_by_variable_uuid_.end() && !c->second.empty())
    throw std::logic_error("Attempting to remove a variable (" + uuid::to_string(u) + ") that is used by existing constraints.");
  variables_.erase(it);
  if (c != constraints_by_variable_uuid_.end()) constraints_by_variable_uuid_.erase(c);
  variables_on_hold_.erase(u);
  ++structure_version_;
  return true;
}
Variable& HashGraph::getVariable(const UUID& u) const {
  auto it = variables_.find(u);
  if (it == variables_.end()) throw std::out_of_range("The variable UUID " + uuid::to_string(u) + " does not exist.");
  return *it->second;
}
void HashGraph::holdVariable(const UUID& u, bool hold) {
  if (hold) variables_on_hold_.insert(u); else variables_on_hold_.erase(u);
  ++structure_version_;
}

}  // namespace bsfuse
