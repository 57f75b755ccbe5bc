// Host-side mirror of the slice of the fuse / beam_slam object model that the solve path reads
// (SURVEY.md section 8b): fuse_core::Variable / Constraint / Loss, the concrete 3-D variable and constraint
// types beam_slam instantiates, and fuse_graphs::HashGraph's containers. Same names, same argument meaning,
// same error behaviour (std::logic_error / std::out_of_range as hash_graph.cpp:128-132,168,233-251), so that
// GpuHashGraph (gpu_hash_graph.h) reads like the subclass a maintainer would add inside fuse_graphs.
//
// Where fuse/ROS headers exist, the same flattener is compiled against the real classes;
// this mirror exists because neither ROS, Boost nor Ceres is installed in this image.
#pragma once
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "uuid.h"

namespace bsfuse {

struct Time {  // ros::Time
  uint32_t sec = 0, nsec = 0;
  Time() = default;
  Time(uint32_t s, uint32_t ns) : sec(s), nsec(ns) {}
  static Time fromSec(double t) {
    Time r; r.sec = static_cast<uint32_t>(t); r.nsec = static_cast<uint32_t>((t - r.sec) * 1e9 + 0.5);
    if (r.nsec >= 1000000000u) { r.sec += 1; r.nsec -= 1000000000u; }
    return r;
  }
  double toSec() const { return sec + 1e-9 * nsec; }
};

// ---- fuse_core::Loss (fuse_core/include/fuse_core/loss.h); plugins in fuse_loss/src ----
struct Loss {
  enum Type { kTrivial = 0, kCauchy = 1, kHuber = 2, kSoftLOne = 3, kArctan = 4, kTukey = 5, kDCS = 6, kFair = 7, kGemanMcClure = 8, kWelsch = 9 };   // = BSNLS_LOSS_*
  Type type = kTrivial;
  double a = 0.0;
  using SharedPtr = std::shared_ptr<Loss>;
};
inline Loss::SharedPtr CauchyLoss(double a) { auto l = std::make_shared<Loss>(); l->type = Loss::kCauchy; l->a = a; return l; }
inline Loss::SharedPtr HuberLoss(double a) { auto l = std::make_shared<Loss>(); l->type = Loss::kHuber; l->a = a; return l; }
// the other one-parameter plugins of fuse_loss (src/*_loss.cpp): SoftLOne, Arctan, Tukey, DCS, Fair, GemanMcClure, Welsch
inline Loss::SharedPtr MakeLoss(Loss::Type type, double a) { auto l = std::make_shared<Loss>(); l->type = type; l->a = a; return l; }

// ---- fuse_core::Variable (fuse_core/include/fuse_core/variable.h:249-352) ----
class Variable {
 public:
  using SharedPtr = std::shared_ptr<Variable>;
  enum Manifold { kEuclidean = 0, kQuaternion = 1 };
  virtual ~Variable() = default;
  const UUID& uuid() const { return uuid_; }
  const std::string& type() const { return type_; }
  size_t size() const { return data_.size(); }
  size_t localSize() const { return manifold_ == kQuaternion ? 3 : data_.size(); }
  double* data() { return data_.data(); }
  const double* data() const { return data_.data(); }
  Manifold manifold() const { return manifold_; }           // stands for localParameterization()
  virtual bool holdConstant() const { return hold_constant_; }
  double lowerBound(size_t) const { return std::numeric_limits<double>::lowest(); }
  double upperBound(size_t) const { return std::numeric_limits<double>::max(); }
  SharedPtr clone() const { return std::make_shared<Variable>(*this); }
  Variable(UUID u, std::string type, std::vector<double> d, Manifold m, bool hold = false)
      : uuid_(u), type_(std::move(type)), data_(std::move(d)), manifold_(m), hold_constant_(hold) {}
  // position of this variable in the last flattened problem of the graph that holds it (GpuHashGraph keeps it; inside
  // fuse the subclass keeps the same number in a side table keyed by the variable pointer)
  mutable int32_t flat_index = -1;
 private:
  UUID uuid_; std::string type_; std::vector<double> data_; Manifold manifold_; bool hold_constant_;
};

// type() strings are boost::typeindex pretty names (variable.h:120)
namespace variables {
inline Variable::SharedPtr Stamped(const char* type, int n, Variable::Manifold m, const Time& t, const UUID& dev) {
  std::vector<double> d(n, 0.0); if (m == Variable::kQuaternion) d[0] = 1.0;
  return std::make_shared<Variable>(uuid::generate(type, t.sec, t.nsec, dev), type, d, m);
}
inline Variable::SharedPtr Orientation3DStamped(const Time& t, const UUID& dev = uuid::nil()) {
  return Stamped("fuse_variables::Orientation3DStamped", 4, Variable::kQuaternion, t, dev); }
inline Variable::SharedPtr Position3DStamped(const Time& t, const UUID& dev = uuid::nil()) {
  return Stamped("fuse_variables::Position3DStamped", 3, Variable::kEuclidean, t, dev); }
inline Variable::SharedPtr VelocityLinear3DStamped(const Time& t, const UUID& dev = uuid::nil()) {
  return Stamped("fuse_variables::VelocityLinear3DStamped", 3, Variable::kEuclidean, t, dev); }
inline Variable::SharedPtr GyroscopeBias3DStamped(const Time& t, const UUID& dev = uuid::nil()) {
  return Stamped("bs_variables::GyroscopeBias3DStamped", 3, Variable::kEuclidean, t, dev); }
inline Variable::SharedPtr AccelerationBias3DStamped(const Time& t, const UUID& dev = uuid::nil()) {
  return Stamped("bs_variables::AccelerationBias3DStamped", 3, Variable::kEuclidean, t, dev); }
inline Variable::SharedPtr Point3DLandmark(uint64_t id) {  // bs_variables/src/point_3d_landmark.cpp:49
  const char* type = "bs_variables::Point3DLandmark";
  return std::make_shared<Variable>(uuid::generate(type, id), type, std::vector<double>(3, 0.0), Variable::kEuclidean);
}
inline Variable::SharedPtr Position3D(const std::string& child, const std::string& parent) {  // position_3d.cpp:16-18
  const char* type = "bs_variables::Position3D";
  return std::make_shared<Variable>(uuid::generate(type, child + parent), type, std::vector<double>(3, 0.0), Variable::kEuclidean);
}
inline Variable::SharedPtr Orientation3D(const std::string& child, const std::string& parent) {  // orientation_3d.cpp:16-18,39-41
  const char* type = "bs_variables::Orientation3D";
  return std::make_shared<Variable>(uuid::generate(type, child + parent), type, std::vector<double>{1, 0, 0, 0},
                                    Variable::kQuaternion, /*holdConstant=*/true);
}
}  // namespace variables

// ---- fuse_core::Constraint (fuse_core/include/fuse_core/constraint.h) ----
// The payload is what each concrete class keeps as members and what its costFunction() hands to Ceres.
class Constraint {
 public:
  using SharedPtr = std::shared_ptr<Constraint>;
  const UUID& uuid() const { return uuid_; }
  const std::string& type() const { return type_; }
  const std::string& source() const { return source_; }
  const std::vector<UUID>& variables() const { return variables_; }
  Loss::SharedPtr loss() const { return loss_; }
  void loss(Loss::SharedPtr l) { loss_ = std::move(l); flat_hint.owner = nullptr; }
  // What the flattener resolved for this constraint the first time it saw it in a graph (type dispatch, variable
  // objects, camera class): the SoA mirror kept up to date at addConstraint time instead of re-deriving it on every
  // optimize (hash_graph.cpp:506-521 re-news every cost function per solve).
  // slots: entries per factor in the group's block list (a marginal prior pads its unused ones with -1); res_index: the
  // constraint's createProblem number in the last flatten (what bsnls_linearize takes)
  struct FlatHint { const void* owner = nullptr; int type = -1, nblocks = 0, slots = 0, nres = 0, stride = 0, camera = -1, res_index = -1; Variable* vars[10] = {}; };
  mutable FlatHint flat_hint;
  const std::vector<double>& payload() const { return payload_; }
  Constraint(UUID u, std::string type, std::string source, std::vector<UUID> vars, std::vector<double> payload)
      : uuid_(u), type_(std::move(type)), source_(std::move(source)), variables_(std::move(vars)),
        payload_(std::move(payload)) {}
 private:
  UUID uuid_; std::string type_, source_; std::vector<UUID> variables_; Loss::SharedPtr loss_;
  std::vector<double> payload_;
};

// constraint uuid source: fuse_core::uuid::generate() (constraint.cpp:46-51). One process-wide generator, seedable.
UUID NextConstraintUuid();
void SeedConstraintUuids(uint64_t seed);

// Dense helpers used by the constraint constructors (sqrt information = cov.inverse().llt().matrixU()).
bool SqrtInformationFromCovariance(const double* cov, int n, double* A /*row-major upper*/);

namespace constraints {
// bs_constraints::EuclideanReprojectionConstraint (euclidean_reprojection_constraint.cpp:17-30).
// payload: pixel[2], K(fx,fy,cx,cy), T_cam_baselink 3x4 row-major [12], weight
Constraint::SharedPtr EuclideanReprojection(const std::string& source, const Variable& R_WORLD_BASELINK,
    const Variable& t_WORLD_BASELINK, const Variable& P_WORLD, const double T_cam_baselink[16],
    const double K[9], const double pixel[2], double reprojection_information_weight);
// bs_constraints::RelativeImuState3DStampedConstraint (relative_imu_state_3d_stamped_constraint.cpp:12-60).
// payload = the BSNLS_IMU_DELTA_STRIDE layout of include/bsnls.h
struct PreintegrationResult;
Constraint::SharedPtr RelativeImuState3DStamped(const std::string& source, const Variable* const state_i[5],
    const Variable* const state_j[5], const double payload288[288]);
// bs_constraints::AbsoluteImuState3DStampedConstraint (absolute_imu_state_3d_stamped_constraint.cpp:13-22)
Constraint::SharedPtr AbsoluteImuState3DStamped(const std::string& source, const Variable* const state[5],
    const double mean[16], const double cov[225]);
// fuse_constraints::RelativePose3DStampedConstraint (relative_pose_3d_stamped_constraint.cpp:46-60)
Constraint::SharedPtr RelativePose3DStamped(const std::string& source, const Variable& p1, const Variable& q1,
    const Variable& p2, const Variable& q2, const double delta[7], const double cov[36]);
// bs_constraints::RelativePose3DStampedWithExtrinsicsConstraint (..._with_extrinsics_constraint.cpp:17-30)
Constraint::SharedPtr RelativePose3DStampedWithExtrinsics(const std::string& source, const Variable& p1,
    const Variable& q1, const Variable& p2, const Variable& q2, const Variable& p_ext, const Variable& q_ext,
    const double delta[7], const double cov[36]);
// fuse_constraints::AbsolutePose3DStampedConstraint (absolute_pose_3d_stamped_constraint.cpp)
Constraint::SharedPtr AbsolutePose3DStamped(const std::string& source, const Variable& p, const Variable& q,
    const double mean[7], const double cov[36]);
// fuse_constraints::RelativeOrientation3DStampedConstraint
Constraint::SharedPtr RelativeOrientation3DStamped(const std::string& source, const Variable& q1,
    const Variable& q2, const double delta[4], const double cov[9]);
// fuse_constraints::AbsoluteOrientation3DStampedConstraint (absolute_orientation_3d_stamped_constraint.cpp:49-58)
Constraint::SharedPtr AbsoluteOrientation3DStamped(const std::string& source, const Variable& q, const double mean[4], const double cov[9]);
// bs_constraints::GravityAlignmentStampedConstraint (gravity_alignment_stamped_constraint.cpp:12-18); cov is 2x2
Constraint::SharedPtr GravityAlignmentStamped(const std::string& source, const Variable& q, const double gravity_in_baselink[3], const double cov[4]);
// bs_constraints::AbsoluteConstraint<T> / RelativeConstraint<T> for 3-vectors (global/absolute_constraint.h:14-28)
Constraint::SharedPtr AbsoluteVec3(const std::string& source, const Variable& x, const double mean[3], const double cov[9]);
Constraint::SharedPtr RelativeVec3(const std::string& source, const Variable& x1, const Variable& x2,
    const double delta[3], const double cov[9]);
}  // namespace constraints

// ---- ceres::Solver::Options / Summary as the graph interface exposes them ----
struct SolverOptions {  // the fields fuse_core/src/ceres_options.cpp:75-200 loads and this path consults
  int max_num_iterations = 50;
  double max_solver_time_in_seconds = 1e9;
  double function_tolerance = 1e-6, gradient_tolerance = 1e-10, parameter_tolerance = 1e-8;
  double initial_trust_region_radius = 1e4, max_trust_region_radius = 1e16, min_trust_region_radius = 1e-32;
  double min_relative_decrease = 1e-3, min_lm_diagonal = 1e-6, max_lm_diagonal = 1e32;
  int max_num_consecutive_invalid_steps = 5;
  bool jacobi_scaling = true;
  int linear_solver = 2;  // BSNLS_LINEAR_AUTO
  int num_threads = 1;    // accepted and ignored on the GPU
  bool IsValid(std::string* error) const;  // ceres::Solver::Options::IsValid subset
};
struct IterationSummary {
  int iteration = 0; bool step_is_valid = false, step_is_successful = false; int linear_solver_iterations = 0;
  double cost = 0, cost_change = 0, gradient_max_norm = 0, gradient_norm = 0, step_norm = 0,
         relative_decrease = 0, trust_region_radius = 0;
};
struct SolverSummary {
  enum TerminationType { CONVERGENCE = 0, NO_CONVERGENCE = 1, FAILURE = 2 };
  TerminationType termination_type = FAILURE;
  std::string message;
  double initial_cost = 0, final_cost = 0, fixed_cost = 0, total_time_in_seconds = 0, device_time_in_seconds = 0;
  int num_successful_steps = 0, num_unsuccessful_steps = 0;
  int num_e_blocks = 0, num_f_blocks = 0, reduced_system_dim = 0, linear_solver_used = 0;
  long long reduced_system_nnz_blocks = 0;
  std::vector<IterationSummary> iterations;
  bool IsSolutionUsable() const { return termination_type == CONVERGENCE || termination_type == NO_CONVERGENCE; }
  std::string BriefReport() const;
  std::string FullReport() const;
};

// ---- fuse_graphs::HashGraph containers (hash_graph.h:367-375, hash_graph.cpp:100-301) ----
class HashGraph {
 public:
  using Variables = std::unordered_map<UUID, Variable::SharedPtr, UuidHash>;
  using Constraints = std::unordered_map<UUID, Constraint::SharedPtr, UuidHash>;
  virtual ~HashGraph() = default;
  void clear();
  bool constraintExists(const UUID& u) const { return constraints_.count(u) > 0; }
  bool addConstraint(Constraint::SharedPtr c);      // std::logic_error if a variable is missing
  bool removeConstraint(const UUID& u);
  const Constraint& getConstraint(const UUID& u) const;  // std::out_of_range
  const Constraints& getConstraints() const { return constraints_; }
  std::vector<UUID> getConnectedConstraints(const UUID& variable) const;
  bool variableExists(const UUID& u) const { return variables_.count(u) > 0; }
  bool addVariable(Variable::SharedPtr v);
  bool removeVariable(const UUID& u);               // std::logic_error if still used
  Variable& getVariable(const UUID& u) const;      // std::out_of_range
  const Variables& getVariables() const { return variables_; }
  void holdVariable(const UUID& u, bool hold_constant = true);
  bool isVariableOnHold(const UUID& u) const { return variables_on_hold_.count(u) > 0; }
  // bumped by every call that changes which blocks / factors exist or which blocks are held; values do not count
  uint64_t structureVersion() const { return structure_version_; }
  // the three virtuals of fuse_core::Graph on the hot path (graph.h; hash_graph.h:317,331-333,355-356)
  virtual SolverSummary optimize(const SolverOptions& options = SolverOptions()) = 0;
  virtual SolverSummary optimizeFor(double max_optimization_time_s, const SolverOptions& options = SolverOptions()) = 0;
  virtual bool evaluate(double* cost, std::vector<double>* residuals, std::vector<double>* gradient) const = 0;
  // getCovariance(covariance_requests, covariance_matrices, options, use_tangent_space) (hash_graph.cpp:303-414): one
  // row-major block per requested variable pair; std::out_of_range for an unknown UUID, std::runtime_error when the
  // blocks cannot be computed
  virtual void getCovariance(const std::vector<std::pair<UUID, UUID>>& covariance_requests,
                             std::vector<std::vector<double>>& covariance_matrices, bool use_tangent_space = true) const = 0;
 protected:
  Constraints constraints_;
  Variables variables_;
  std::unordered_set<UUID, UuidHash> variables_on_hold_;
  std::unordered_map<UUID, std::unordered_set<UUID, UuidHash>, UuidHash> constraints_by_variable_uuid_;
  uint64_t structure_version_ = 1;
};

}  // namespace bsfuse
