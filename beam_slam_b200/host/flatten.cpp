#include "flatten.h"

#include <algorithm>
#include <array>
#include <functional>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <mutex>
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
template <typename F> void ForBlocks(size_t n, const F& body, size_t grain = 16384) {
  constexpr size_t kMaxThreads = BSFUSE_COPY_THREADS;
  const size_t nt = std::min<size_t>(std::min<size_t>(kMaxThreads, std::max(1u, std::thread::hardware_concurrency())), n / grain);
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
// First sight of a constraint in this graph, part 1 (thread-safe: reads the graph, writes only the constraint's own hint):
// type() string dispatch and the variable objects. `last` caches the previous constraint's type entry (graphs hold long
// runs of one type).
void ResolveHintTypeAndVariables(const HashGraph& graph, const Constraint& c, const std::pair<const std::string, TypeInfo>** last) {
  if (!*last || (*last)->first != c.type()) {
    auto ti = TypeTable().find(c.type());
    if (ti == TypeTable().end()) throw std::invalid_argument("constraint type not supported by the GPU path: " + c.type());
    *last = &*ti;
  }
  const TypeInfo& info = (*last)->second;
  const int nvar = static_cast<int>(c.variables().size());
  if (info.type == BSNLS_FACTOR_MARGINAL ? (nvar < 1 || nvar > info.nblocks) : nvar != info.nblocks) throw std::invalid_argument("unexpected variable count for " + c.type());
  Constraint::FlatHint& hnt = c.flat_hint;
  hnt.type = info.type; hnt.nblocks = nvar; hnt.slots = info.nblocks; hnt.nres = info.nres; hnt.stride = info.stride; hnt.camera = -1;
  for (int b = 0; b < nvar; ++b) hnt.vars[b] = &graph.getVariable(c.variables()[b]);  // variables_.at(uuid), :515
}
}  // namespace

void Flatten(const HashGraph& graph, FlatProblem* out) {
  const auto T0 = std::chrono::steady_clock::now(); auto lap = [&](const char* n) { if (std::getenv("BSFUSE_DEBUG")) std::fprintf(stderr, "[bsfuse] flatten %-20s %8.3f ms\n", n, 1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - T0).count()); };
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
  lap("variables");
  if (P.groups.size() != BSNLS_NUM_FACTOR_TYPES) P.groups.assign(BSNLS_NUM_FACTOR_TYPES, FlatProblem::Group());
  for (int t = 0; t < BSNLS_NUM_FACTOR_TYPES; ++t) {
    FlatProblem::Group& g = P.groups[t];
    g.type = t; g.blocks.clear(); g.aux.clear(); g.loss_type.clear(); g.res_index.clear(); g.consts.clear(); g.loss_param.clear();
  }
  // createProblem, second loop (hash_graph.cpp:506-521). One sequential walk over the container fixes the order (residual
  // block i = the i-th constraint of the walk); everything per constraint -- the hash lookups of its variables the first
  // time it is seen, and writing its row of the SoA groups -- then runs on chunks of that list in parallel, each chunk
  // at offsets known from a per-type count, so the result does not depend on the thread count.
  std::vector<const Constraint*> all; all.reserve(graph.getConstraints().size());
  for (const auto& kv : graph.getConstraints()) all.push_back(kv.second.get());
  lap("constraint walk");
  const size_t nc = all.size();
  constexpr size_t kChunk = 16384;
  const size_t nchunks = std::max<size_t>(1, (nc + kChunk - 1) / kChunk);
  std::exception_ptr err; std::mutex err_mutex;
  auto guarded = [&](const std::function<void(size_t)>& body) {     // body(chunk) on every chunk; the first exception is re-thrown
    ForBlocks(nchunks, [&](size_t c0, size_t c1) {
      try { for (size_t ch = c0; ch < c1; ++ch) body(ch); }
      catch (...) { std::lock_guard<std::mutex> lock(err_mutex); if (!err) err = std::current_exception(); }
    }, 1);
    if (err) std::rethrow_exception(err);
  };
  // (1) constraints not seen before: type dispatch + variable objects; the camera classes each chunk meets, in order
  std::vector<std::vector<std::array<double, BSNLS_CAMERA_STRIDE>>> chunk_cams(nchunks);
  auto camera_key = [](const Constraint& c, double* key) {
    const auto& pl = c.payload();
    std::memcpy(key, pl.data() + 2, sizeof(double) * 17);  // K(4), T(12), w
    key[17] = c.loss() ? static_cast<double>(c.loss()->type) : static_cast<double>(BSNLS_LOSS_NONE);
    key[18] = c.loss() ? c.loss()->a : 0.0; key[19] = 0.0;
  };
  guarded([&](size_t ch) {
    const std::pair<const std::string, TypeInfo>* last = nullptr;
    std::array<double, BSNLS_CAMERA_STRIDE> key, prev; bool have_prev = false;
    for (size_t i = ch * kChunk; i < std::min(nc, (ch + 1) * kChunk); ++i) {
      const Constraint& c = *all[i];
      if (c.flat_hint.owner == &P) continue;
      ResolveHintTypeAndVariables(graph, c, &last);
      if (c.flat_hint.type != BSNLS_FACTOR_REPROJECTION) continue;
      camera_key(c, key.data());
      if (have_prev && key == prev) continue;
      prev = key; have_prev = true;
      auto& seen = chunk_cams[ch];
      if (std::find(seen.begin(), seen.end(), key) == seen.end()) seen.push_back(key);
    }
  });
  for (size_t ch = 0; ch < nchunks; ++ch)            // camera table in order of first appearance along the walk
    for (const auto& key : chunk_cams[ch]) {
      std::vector<double> cam(key.begin(), key.end());
      if (P.camera_index.find(cam) == P.camera_index.end()) {
        P.camera_index.emplace(cam, static_cast<int32_t>(P.cameras.size() / BSNLS_CAMERA_STRIDE));
        P.cameras.insert(P.cameras.end(), cam.begin(), cam.end());
      }
    }
  lap("new constraints");
  // (2) per-type counts of every chunk -> where each chunk writes
  std::vector<std::array<int64_t, BSNLS_NUM_FACTOR_TYPES>> count(nchunks + 1);
  guarded([&](size_t ch) {
    count[ch + 1].fill(0);
    for (size_t i = ch * kChunk; i < std::min(nc, (ch + 1) * kChunk); ++i) count[ch + 1][all[i]->flat_hint.type]++;
  });
  count[0].fill(0);
  for (size_t ch = 0; ch < nchunks; ++ch) for (int t = 0; t < BSNLS_NUM_FACTOR_TYPES; ++t) count[ch + 1][t] += count[ch][t];
  static const int kSlots[BSNLS_NUM_FACTOR_TYPES] = {3, 10, 5, 4, 6, 2, 2, 1, 2, BSNLS_MARGINAL_MAX_BLOCKS, 1, 1};
  static const int kStrides[BSNLS_NUM_FACTOR_TYPES] = {BSNLS_REPROJECTION_STRIDE, BSNLS_IMU_DELTA_STRIDE, BSNLS_IMU_PRIOR_STRIDE, BSNLS_POSE_STRIDE, BSNLS_POSE_STRIDE,
      BSNLS_POSE_STRIDE, BSNLS_ORIENTATION_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_MARGINAL_STRIDE, BSNLS_GRAVITY_STRIDE, BSNLS_ORIENTATION_STRIDE};
  for (int t = 0; t < BSNLS_NUM_FACTOR_TYPES; ++t) {
    FlatProblem::Group& g = P.groups[t];
    const size_t n = static_cast<size_t>(count[nchunks][t]);
    g.blocks.assign(n * kSlots[t], -1); g.res_index.assign(n, 0); g.consts.assign(n * kStrides[t], 0.0);
    if (t == BSNLS_FACTOR_REPROJECTION) g.aux.assign(n, 0); else { g.loss_type.assign(n, BSNLS_LOSS_NONE); g.loss_param.assign(n, 0.0); }
  }
  // (3) the rows
  guarded([&](size_t ch) {
    std::array<int64_t, BSNLS_NUM_FACTOR_TYPES> at = count[ch];
    std::array<double, BSNLS_CAMERA_STRIDE> key, prev; int32_t prev_cam = -1;
    for (size_t i = ch * kChunk; i < std::min(nc, (ch + 1) * kChunk); ++i) {
      const Constraint& c = *all[i];
      Constraint::FlatHint& hnt = c.flat_hint;
      FlatProblem::Group& g = P.groups[hnt.type];
      const size_t k = static_cast<size_t>(at[hnt.type]++);
      for (int b = 0; b < hnt.nblocks; ++b) g.blocks[k * hnt.slots + b] = hnt.vars[b]->flat_index;
      const auto& pl = c.payload();
      if (hnt.type == BSNLS_FACTOR_REPROJECTION) {
        g.consts[2 * k] = pl[0]; g.consts[2 * k + 1] = pl[1];
        if (hnt.owner != &P) {                       // camera class: the table is complete, lookups only
          camera_key(c, key.data());
          if (prev_cam < 0 || !(key == prev)) { prev_cam = P.camera_index.at(std::vector<double>(key.begin(), key.end())); prev = key; }
          hnt.camera = prev_cam;
        }
        g.aux[k] = hnt.camera;
      } else {
        std::memcpy(g.consts.data() + k * hnt.stride, pl.data(), sizeof(double) * std::min<size_t>(pl.size(), hnt.stride));
        g.loss_type[k] = c.loss() ? static_cast<int>(c.loss()->type) : BSNLS_LOSS_NONE;
        g.loss_param[k] = c.loss() ? c.loss()->a : 0.0;
      }
      hnt.owner = &P;   // variable objects belong to the graph, the camera index to this FlatProblem's table: one FlatProblem serves one graph
      hnt.res_index = static_cast<int32_t>(i);
      g.res_index[k] = static_cast<int32_t>(i);
    }
  });
  lap("constraints");
  P.Finalize();
}

}  // namespace bsfuse
