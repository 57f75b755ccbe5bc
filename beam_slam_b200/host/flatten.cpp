#include "flatten.h"

#include <algorithm>
#include <cstring>
#include <thread>

namespace bsfuse {

namespace {
struct TypeInfo { int type; int nblocks; int stride; int nres; };
const std::unordered_map<std::string, TypeInfo>& TypeTable() {
  static const std::unordered_map<std::string, TypeInfo> t = {
      {"bs_constraints::EuclideanReprojectionConstraint", {BSNLS_FACTOR_REPROJECTION, 3, BSNLS_REPROJECTION_STRIDE, 2}},
      {"bs_constraints::RelativeImuState3DStampedConstraint", {BSNLS_FACTOR_IMU_DELTA, 10, BSNLS_IMU_DELTA_STRIDE, 15}},
      {"bs_constraints::AbsoluteImuState3DStampedConstraint", {BSNLS_FACTOR_IMU_PRIOR, 5, BSNLS_IMU_PRIOR_STRIDE, 15}},
      {"fuse_constraints::RelativePose3DStampedConstraint", {BSNLS_FACTOR_RELATIVE_POSE, 4, BSNLS_POSE_STRIDE, 6}},
      {"bs_constraints::RelativePose3DStampedWithExtrinsicsConstraint", {BSNLS_FACTOR_RELATIVE_POSE_EXTRINSICS, 6, BSNLS_POSE_STRIDE, 6}},
      {"fuse_constraints::AbsolutePose3DStampedConstraint", {BSNLS_FACTOR_ABSOLUTE_POSE, 2, BSNLS_POSE_STRIDE, 6}},
      {"fuse_constraints::RelativeOrientation3DStampedConstraint", {BSNLS_FACTOR_RELATIVE_ORIENTATION, 2, BSNLS_ORIENTATION_STRIDE, 3}},
      {"bs_constraints::AbsoluteConstraint3D", {BSNLS_FACTOR_ABSOLUTE_VEC3, 1, BSNLS_VEC3_STRIDE, 3}},
      {"bs_constraints::RelativeConstraint3D", {BSNLS_FACTOR_RELATIVE_VEC3, 2, BSNLS_VEC3_STRIDE, 3}},
      {"bs_constraints::AbsolutePose3DConstraint", {BSNLS_FACTOR_ABSOLUTE_POSE, 2, BSNLS_POSE_STRIDE, 6}},   // same functor as fuse's (absolute_pose_3d_constraint.cpp:45-50)
      {"bs_constraints::GravityAlignmentStampedConstraint", {BSNLS_FACTOR_GRAVITY_ALIGNMENT, 1, BSNLS_GRAVITY_STRIDE, 2}},
      {"fuse_constraints::AbsoluteOrientation3DStampedConstraint", {BSNLS_FACTOR_ABSOLUTE_ORIENTATION, 1, BSNLS_ORIENTATION_STRIDE, 3}},
      // the only type with a variable number of blocks: nblocks here = slots in the block list (marginal_constraint.h:85-96)
      {"fuse_constraints::MarginalConstraint", {BSNLS_FACTOR_MARGINAL, BSNLS_MARGINAL_MAX_BLOCKS, BSNLS_MARGINAL_STRIDE, BSNLS_MARGINAL_MAX_ROWS}},
  };
  return t;
}
}  // namespace

void FlatProblem::Finalize() {
  cgroups.clear();
  for (auto& g : groups) {
    if (g.res_index.empty()) continue;
    bsnls_factor_group cg{};
    cg.type = g.type; cg.count = static_cast<int32_t>(g.res_index.size());
    cg.blocks = g.blocks.data(); cg.consts = g.consts.data();
    cg.aux = g.aux.empty() ? nullptr : g.aux.data();
    cg.loss_type = g.loss_type.empty() ? nullptr : g.loss_type.data();
    cg.loss_param = g.loss_param.empty() ? nullptr : g.loss_param.data();
    cg.res_index = g.res_index.data();
    cgroups.push_back(cg);
  }
  c.num_blocks = static_cast<int32_t>(block_size.size());
  c.block_size = block_size.data(); c.block_manifold = block_manifold.data();
  c.block_constant = block_constant.data(); c.block_group = block_group.data();
  c.block_values = block_values.data();
  c.num_cameras = static_cast<int32_t>(cameras.size() / BSNLS_CAMERA_STRIDE);
  c.cameras = cameras.data();
  c.num_groups = static_cast<int32_t>(cgroups.size());
  c.groups = cgroups.data();
}

namespace {
// Variable values live in one heap object per variable (as in fuse): copying 200 k of them one by one is bound by cache
// misses, so large graphs split the range over a few threads (a window of some hundred variables stays on the caller's).
#ifndef BSFUSE_COPY_THREADS
#define BSFUSE_COPY_THREADS 8
#endif
template <typename F> void ForBlocks(size_t n, const F& body) {
  constexpr size_t kGrain = 16384, kMaxThreads = BSFUSE_COPY_THREADS;
  const size_t nt = std::min<size_t>(std::min<size_t>(kMaxThreads, std::max(1u, std::thread::hardware_concurrency())), n / kGrain);
  if (nt <= 1) { body(0, n); return; }
  std::vector<std::thread> pool;
  for (size_t t = 1; t < nt; ++t) pool.emplace_back([&, t] { body(n * t / nt, n * (t + 1) / nt); });
  body(0, n / nt);
  for (auto& th : pool) th.join();
}
}  // namespace

void FlatProblem::ScatterValues(const double* values) const {
  ForBlocks(block_variable.size(), [&](size_t b0, size_t b1) {
    for (size_t b = b0; b < b1; ++b) std::memcpy(block_variable[b]->data(), values + block_offset[b], sizeof(double) * block_size[b]);
  });
}

void FlatProblem::GatherValues() {
  ForBlocks(block_variable.size(), [&](size_t b0, size_t b1) {
    for (size_t b = b0; b < b1; ++b) std::memcpy(block_values.data() + block_offset[b], block_variable[b]->data(), sizeof(double) * block_size[b]);
  });
}

namespace {
// first sight of a constraint in this graph: type() string dispatch, variable objects, camera class
void ResolveHint(const HashGraph& graph, const Constraint& c, FlatProblem& P) {
  auto ti = TypeTable().find(c.type());
  if (ti == TypeTable().end()) throw std::invalid_argument("constraint type not supported by the GPU path: " + c.type());
  const TypeInfo& info = ti->second;
  const int nvar = static_cast<int>(c.variables().size());
  if (info.type == BSNLS_FACTOR_MARGINAL ? (nvar < 1 || nvar > info.nblocks) : nvar != info.nblocks) throw std::invalid_argument("unexpected variable count for " + c.type());
  Constraint::FlatHint& hnt = c.flat_hint;
  hnt.type = info.type; hnt.nblocks = nvar; hnt.slots = info.nblocks; hnt.nres = info.nres; hnt.stride = info.stride; hnt.camera = -1;
  for (int b = 0; b < nvar; ++b) hnt.vars[b] = &graph.getVariable(c.variables()[b]);  // variables_.at(uuid), :515
  if (info.type == BSNLS_FACTOR_REPROJECTION) {
    const auto& pl = c.payload();
    std::vector<double> cam(pl.begin() + 2, pl.begin() + 19);  // K(4), T(12), w
    cam.push_back(c.loss() ? static_cast<double>(c.loss()->type) : static_cast<double>(BSNLS_LOSS_NONE));
    cam.push_back(c.loss() ? c.loss()->a : 0.0); cam.push_back(0.0);
    auto it = P.camera_index.find(cam);
    if (it == P.camera_index.end()) {
      it = P.camera_index.emplace(cam, static_cast<int32_t>(P.cameras.size() / BSNLS_CAMERA_STRIDE)).first;
      P.cameras.insert(P.cameras.end(), cam.begin(), cam.end());
    }
    hnt.camera = it->second;
  }
  hnt.owner = &P;   // variable objects belong to the graph, the camera index to this FlatProblem's table: one FlatProblem serves one graph
}
}  // namespace

void Flatten(const HashGraph& graph, FlatProblem* out) {
  FlatProblem& P = *out;
  const auto& vars = graph.getVariables();
  P.block_size.clear(); P.block_manifold.clear(); P.block_group.clear(); P.block_constant.clear(); P.block_values.clear();
  P.block_offset.clear(); P.block_variable.clear(); P.block_uuid.clear();
  P.block_size.reserve(vars.size()); P.block_values.reserve(vars.size() * 4);
  int64_t off = 0;
  // createProblem, first loop (hash_graph.cpp:478-504)
  for (const auto& kv : vars) {
    Variable& v = *kv.second;
    for (size_t i = 0; i < v.size(); ++i)
      if (v.lowerBound(i) > std::numeric_limits<double>::lowest() || v.upperBound(i) < std::numeric_limits<double>::max())
        throw std::invalid_argument("bounded variables are not supported by the GPU path: " + v.type());
    v.flat_index = static_cast<int32_t>(P.block_size.size());
    P.block_size.push_back(static_cast<int32_t>(v.size()));
    P.block_manifold.push_back(v.manifold() == Variable::kQuaternion ? BSNLS_MANIFOLD_QUATERNION : BSNLS_MANIFOLD_EUCLIDEAN);
    P.block_constant.push_back((v.holdConstant() || graph.isVariableOnHold(kv.first)) ? 1 : 0);
    // elimination groups: landmarks first (explicit two-group ordering)
    const std::string& ty = v.type();
    const bool landmark = ty.size() > 14 && ty.compare(0, 14, "bs_variables::") == 0 && (ty == "bs_variables::Point3DLandmark" || ty == "bs_variables::InverseDepthLandmark");
    P.block_group.push_back(landmark ? 0 : 1);
    P.block_offset.push_back(off);
    P.block_variable.push_back(&v);
    P.block_uuid.push_back(kv.first);
    for (size_t i = 0; i < v.size(); ++i) P.block_values.push_back(v.data()[i]);
    off += static_cast<int64_t>(v.size());
  }
  if (P.groups.size() != BSNLS_NUM_FACTOR_TYPES) P.groups.assign(BSNLS_NUM_FACTOR_TYPES, FlatProblem::Group());
  for (int t = 0; t < BSNLS_NUM_FACTOR_TYPES; ++t) {
    FlatProblem::Group& g = P.groups[t];
    g.type = t; g.blocks.clear(); g.aux.clear(); g.loss_type.clear(); g.res_index.clear(); g.consts.clear(); g.loss_param.clear();
  }
  int32_t res_index = 0;
  // createProblem, second loop (hash_graph.cpp:506-521)
  for (const auto& kv : graph.getConstraints()) {
    const Constraint& c = *kv.second;
    if (c.flat_hint.owner != &P) ResolveHint(graph, c, P);
    const Constraint::FlatHint& hnt = c.flat_hint;
    FlatProblem::Group& g = P.groups[hnt.type];
    for (int b = 0; b < hnt.slots; ++b) g.blocks.push_back(b < hnt.nblocks ? hnt.vars[b]->flat_index : -1);
    const auto& pl = c.payload();
    if (hnt.type == BSNLS_FACTOR_REPROJECTION) {
      g.consts.push_back(pl[0]); g.consts.push_back(pl[1]);
      g.aux.push_back(hnt.camera);
    } else {
      const size_t n0 = g.consts.size();
      g.consts.resize(n0 + hnt.stride, 0.0);
      std::memcpy(g.consts.data() + n0, pl.data(), sizeof(double) * std::min<size_t>(pl.size(), hnt.stride));
      g.loss_type.push_back(c.loss() ? static_cast<int>(c.loss()->type) : BSNLS_LOSS_NONE);
      g.loss_param.push_back(c.loss() ? c.loss()->a : 0.0);
    }
    c.flat_hint.res_index = res_index;
    g.res_index.push_back(res_index++);
  }
  P.Finalize();
}

}  // namespace bsfuse
