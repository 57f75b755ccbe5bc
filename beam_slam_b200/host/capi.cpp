// C interface over the host-side mirror, for the ctypes plumbing in beam_slam_b200/host.py.
#include <cstring>
#include <cmath>
#include <memory>
#include <random>

#include "gpu_hash_graph.h"
#include "marginalize.h"
#include "preintegrator.h"
#include "smoother_replay.h"
#include "synthetic.h"

using namespace bsfuse;

struct bsfuse_graph {
  GpuHashGraph graph;
  const FlatProblem& flat() const { return graph.lastProblem(); }   // the graph's own SoA view (bsfuse_flatten refreshes it)
  std::vector<std::pair<UUID, std::vector<double>>> truth;
  SolverSummary last;
  std::string error;
};

extern "C" {

bsfuse_graph* bsfuse_graph_new() { return new bsfuse_graph(); }
void bsfuse_graph_free(bsfuse_graph* g) { delete g; }
const char* bsfuse_last_error(bsfuse_graph* g) { return g->error.c_str(); }

int bsfuse_build_synthetic(bsfuse_graph* g, int config, double scale, uint64_t seed, long long info_out[8]) {
  try {
    g->truth.clear();
    const SyntheticInfo i = BuildSynthetic(config, scale, seed, &g->graph, &g->truth);
    const long long v[8] = {i.num_keyframes, i.num_landmarks, i.num_observations, i.num_imu, i.num_relative_pose,
                            i.num_priors, i.num_variables, i.num_constraints};
    std::memcpy(info_out, v, sizeof(v));
    return 0;
  } catch (const std::exception& e) { g->error = e.what(); return -1; }
}
void bsfuse_synthetic_options(int config, bsnls_options* out) { ToCOptions(SyntheticSolverOptions(config), out); }

// createProblem-order flattening; the returned struct and its arrays stay valid until the next flatten/free.
const bsnls_problem* bsfuse_flatten(bsfuse_graph* g) {
  try { return &g->graph.flattened().c; }
  catch (const std::exception& e) { g->error = e.what(); return nullptr; }
}
long long bsfuse_num_values(bsfuse_graph* g) { return static_cast<long long>(g->flat().block_values.size()); }
void bsfuse_block_uuid(bsfuse_graph* g, int block, unsigned char out[16]) { std::memcpy(out, g->flat().block_uuid[block].data.data(), 16); }
// ground truth in the flattened block order (NaN where unknown)
void bsfuse_truth_values(bsfuse_graph* g, double* out) {
  std::unordered_map<UUID, const std::vector<double>*, UuidHash> m;
  for (const auto& kv : g->truth) m[kv.first] = &kv.second;
  for (size_t b = 0; b < g->flat().block_uuid.size(); ++b) {
    auto it = m.find(g->flat().block_uuid[b]);
    for (int k = 0; k < g->flat().block_size[b]; ++k)
      out[g->flat().block_offset[b] + k] = it == m.end() ? std::nan("") : (*it->second)[k];
  }
}
// current Variable::data() in the flattened block order
void bsfuse_current_values(bsfuse_graph* g, double* out) {
  for (size_t b = 0; b < g->flat().block_variable.size(); ++b)
    std::memcpy(out + g->flat().block_offset[b], g->flat().block_variable[b]->data(), sizeof(double) * g->flat().block_size[b]);
}
void bsfuse_set_values(bsfuse_graph* g, const double* values) { g->flat().ScatterValues(values); }

}  // extern "C"

namespace {
SolverOptions FromC(const bsnls_options* o) {
  SolverOptions so;
  so.max_num_iterations = o->max_num_iterations; so.max_solver_time_in_seconds = o->max_solver_time_in_seconds;
  so.function_tolerance = o->function_tolerance; so.gradient_tolerance = o->gradient_tolerance;
  so.parameter_tolerance = o->parameter_tolerance; so.initial_trust_region_radius = o->initial_trust_region_radius;
  so.max_trust_region_radius = o->max_trust_region_radius; so.min_trust_region_radius = o->min_trust_region_radius;
  so.min_relative_decrease = o->min_relative_decrease; so.min_lm_diagonal = o->min_lm_diagonal;
  so.max_lm_diagonal = o->max_lm_diagonal; so.max_num_consecutive_invalid_steps = o->max_num_consecutive_invalid_steps;
  so.jacobi_scaling = o->jacobi_scaling != 0; so.linear_solver = o->linear_solver;
  return so;
}
void ToC(const SolverSummary& r, bsnls_summary* s, bsnls_iteration* its, int max_its) {
  std::memset(s, 0, sizeof(*s));
  s->termination_type = r.termination_type; s->num_iterations = static_cast<int>(r.iterations.size());
  s->num_successful_steps = r.num_successful_steps; s->num_unsuccessful_steps = r.num_unsuccessful_steps;
  s->initial_cost = r.initial_cost; s->final_cost = r.final_cost; s->fixed_cost = r.fixed_cost;
  s->total_time_in_seconds = r.total_time_in_seconds; s->device_time_in_seconds = r.device_time_in_seconds;
  s->num_e_blocks = r.num_e_blocks; s->num_f_blocks = r.num_f_blocks; s->reduced_system_dim = r.reduced_system_dim;
  s->linear_solver_used = r.linear_solver_used; s->reduced_system_nnz_blocks = r.reduced_system_nnz_blocks;
  std::snprintf(s->message, sizeof(s->message), "%s", r.message.c_str());
  for (int i = 0; i < s->num_iterations && i < max_its && its; ++i) {
    const IterationSummary& a = r.iterations[i];
    its[i] = bsnls_iteration{a.iteration, a.step_is_valid, a.step_is_successful, a.linear_solver_iterations, a.cost,
                             a.cost_change, a.gradient_max_norm, a.gradient_norm, a.step_norm, a.relative_decrease,
                             a.trust_region_radius};
  }
}
void TimingOut(const GpuHashGraph& g, double out[5]) {
  const auto& t = g.lastTiming();
  out[0] = t.flatten; out[1] = t.upload; out[2] = t.solve; out[3] = t.download; out[4] = t.structure_reused ? 1.0 : 0.0;
}
}  // namespace

struct bsfuse_replay { std::unique_ptr<SmootherReplay> r; std::string error; };

extern "C" {

// GpuHashGraph::optimize / optimizeFor (max_time_s >= 0) through the public interface; fills the C summary for python.
int bsfuse_optimize_for(bsfuse_graph* g, double max_time_s, const bsnls_options* o, bsnls_summary* s, bsnls_iteration* its, int max_its) {
  try {
    const SolverOptions so = FromC(o);
    g->last = max_time_s >= 0 ? g->graph.optimizeFor(max_time_s, so) : g->graph.optimize(so);
    ToC(g->last, s, its, max_its);
    return 0;
  } catch (const std::invalid_argument& e) { g->error = e.what(); return -1; }
  catch (const std::exception& e) { g->error = e.what(); return -3; }
}
int bsfuse_optimize(bsfuse_graph* g, const bsnls_options* o, bsnls_summary* s, bsnls_iteration* its, int max_its) {
  return bsfuse_optimize_for(g, -1.0, o, s, its, max_its);
}
// host wall-clock split of the last optimize: flatten | set_problem or set_values | solve | get_values + scatter | reused (0/1)
void bsfuse_last_timing(bsfuse_graph* g, double out[5]) { TimingOut(g->graph, out); }

// GpuHashGraph::getCovariance for pairs given as block indices of the last flatten (mapped to UUIDs here, so the call runs
// through the graph interface); out: 16 doubles reserved per pair
int bsfuse_get_covariance(bsfuse_graph* g, const int* block_pairs, int n, int use_tangent_space, double* out) {
  try {
    const FlatProblem& f = g->graph.flattened();
    std::vector<std::pair<UUID, UUID>> req;
    for (int i = 0; i < n; ++i) req.emplace_back(f.block_uuid.at(block_pairs[2 * i]), f.block_uuid.at(block_pairs[2 * i + 1]));
    std::vector<std::vector<double>> m;
    g->graph.getCovariance(req, m, use_tangent_space != 0);
    for (int i = 0; i < n; ++i) { std::memset(out + 16 * i, 0, sizeof(double) * 16); std::memcpy(out + 16 * i, m[i].data(), sizeof(double) * m[i].size()); }
    return 0;
  } catch (const std::out_of_range& e) { g->error = e.what(); return -1; }
  catch (const std::exception& e) { g->error = e.what(); return -3; }
}

// ---- marginalisation (marginalize.h) ----
// A small pose graph without landmarks for the marginalisation tests: n stamped poses on a helix, a prior on the first,
// odometry between neighbours and the longer links 0-2, 1-3 (and 2-4 from six poses on), measurements from the truth plus noise, values perturbed.
int bsfuse_build_pose_chain(bsfuse_graph* g, int n, uint64_t seed) {
  try {
    if (n < 4) throw std::invalid_argument("pose chain: at least four poses");
    std::mt19937_64 rng(seed); std::normal_distribution<double> N(0.0, 1.0);
    struct Pose { double q[4], p[3]; Variable::SharedPtr vq, vp; };
    std::vector<Pose> ps(n);
    auto qmul = [](const double* a, const double* b, double* r) {
      r[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3]; r[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
      r[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1]; r[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]; };
    auto qrot = [&](const double* q, const double* v, double* r) { const double vq[4] = {0, v[0], v[1], v[2]}, qc[4] = {q[0], -q[1], -q[2], -q[3]}; double t[4], u[4]; qmul(q, vq, t); qmul(t, qc, u); r[0] = u[1]; r[1] = u[2]; r[2] = u[3]; };
    auto small_q = [&](double sigma, double* q) { const double a[3] = {sigma * N(rng), sigma * N(rng), sigma * N(rng)}; const double nrm = std::sqrt(1.0 + 0.25 * (a[0] * a[0] + a[1] * a[1] + a[2] * a[2])); q[0] = 1.0 / nrm; for (int k = 0; k < 3; ++k) q[1 + k] = 0.5 * a[k] / nrm; };
    for (int k = 0; k < n; ++k) {
      const double th = 0.35 * k;
      ps[k].q[0] = std::cos(0.5 * th); ps[k].q[1] = 0.1 * std::sin(0.5 * th); ps[k].q[2] = 0.2 * std::sin(0.5 * th); ps[k].q[3] = std::sqrt(1.0 - 0.05) * std::sin(0.5 * th);
      const double nq = std::sqrt(ps[k].q[0] * ps[k].q[0] + ps[k].q[1] * ps[k].q[1] + ps[k].q[2] * ps[k].q[2] + ps[k].q[3] * ps[k].q[3]);
      for (double& v : ps[k].q) v /= nq;
      ps[k].p[0] = 2.0 * std::cos(th); ps[k].p[1] = 2.0 * std::sin(th); ps[k].p[2] = 0.1 * k;
      const Time t(10 + k, 0);
      ps[k].vq = variables::Orientation3DStamped(t); ps[k].vp = variables::Position3DStamped(t);
      double dq[4], qv[4]; small_q(0.02, dq); qmul(ps[k].q, dq, qv);
      for (int i = 0; i < 4; ++i) ps[k].vq->data()[i] = qv[i];
      for (int i = 0; i < 3; ++i) ps[k].vp->data()[i] = ps[k].p[i] + 0.05 * N(rng);
      g->graph.addVariable(ps[k].vp); g->graph.addVariable(ps[k].vq);
    }
    double cov[36] = {0}; for (int i = 0; i < 6; ++i) cov[i * 6 + i] = i < 3 ? 4e-4 : 1e-4;
    double mean[7] = {ps[0].p[0], ps[0].p[1], ps[0].p[2], ps[0].q[0], ps[0].q[1], ps[0].q[2], ps[0].q[3]};
    g->graph.addConstraint(constraints::AbsolutePose3DStamped("PRIOR", *ps[0].vp, *ps[0].vq, mean, cov));
    auto rel = [&](int a, int b) {
      const double qa_c[4] = {ps[a].q[0], -ps[a].q[1], -ps[a].q[2], -ps[a].q[3]};
      double d[3] = {ps[b].p[0] - ps[a].p[0], ps[b].p[1] - ps[a].p[1], ps[b].p[2] - ps[a].p[2]}, dp[3], dq[4], nq[4], dqn[4];
      qrot(qa_c, d, dp); qmul(qa_c, ps[b].q, dq); small_q(0.01, nq); qmul(dq, nq, dqn);
      const double delta[7] = {dp[0] + 0.02 * N(rng), dp[1] + 0.02 * N(rng), dp[2] + 0.02 * N(rng), dqn[0], dqn[1], dqn[2], dqn[3]};
      g->graph.addConstraint(constraints::RelativePose3DStamped("ODOM", *ps[a].vp, *ps[a].vq, *ps[b].vp, *ps[b].vq, delta, cov));
    };
    for (int k = 0; k + 1 < n; ++k) rel(k, k + 1);
    rel(0, 2); rel(1, 3);
    if (n >= 6) rel(2, 4);   // pose 2 then has four neighbours: its marginal is 18 rows tall
    return 0;
  } catch (const std::exception& e) { g->error = e.what(); return -1; }
}
// The graph of test_marginalize_variables.cpp:416-470 (MarginalizeVariables), literal values: orientations x1, x2, x3, l1
// (stamps 1, 2, 3 and 3.5 s), a prior on x1 and the relative orientations x1-x2, x2-x3, x2-l1, covariance diag(1, 2, 3).
int bsfuse_build_orientation_chain(bsfuse_graph* g) {
  try {
    auto make = [&](const Time& t, double w, double x, double y, double z) {
      auto v = variables::Orientation3DStamped(t); v->data()[0] = w; v->data()[1] = x; v->data()[2] = y; v->data()[3] = z; g->graph.addVariable(v); return v; };
    auto x1 = make(Time(1, 0), 0.927362, 0.1, 0.2, 0.3), x2 = make(Time(2, 0), 0.848625, 0.13798, 0.175959, 0.479411);
    auto x3 = make(Time(3, 0), 0.735597, 0.170384, 0.144808, 0.63945), l1 = make(Time(3, 500000000), 0.803884, 0.304917, 0.268286, 0.434533);
    const double cov[9] = {1, 0, 0, 0, 2, 0, 0, 0, 3};
    const double mean1[4] = {0.92736185, 0.1, 0.2, 0.3}, d23[4] = {0.979795897, 0.0, 0.0, 0.2}, d4[4] = {0.979795897, 0.2, 0.0, 0.0};
    g->graph.addConstraint(constraints::AbsoluteOrientation3DStamped("test", *x1, mean1, cov));
    g->graph.addConstraint(constraints::RelativeOrientation3DStamped("test", *x1, *x2, d23, cov));
    g->graph.addConstraint(constraints::RelativeOrientation3DStamped("test", *x2, *x3, d23, cov));
    g->graph.addConstraint(constraints::RelativeOrientation3DStamped("test", *x2, *l1, d4, cov));
    return 0;
  } catch (const std::exception& e) { g->error = e.what(); return -1; }
}
// marginalizeVariables + Graph::update for the variables with the given UUIDs (16 bytes each). out: constraints removed,
// marginal constraints added, variables removed. -1: unknown variable, -2: a marginal does not fit the GPU factor, -3: other
int bsfuse_marginalize(bsfuse_graph* g, const unsigned char* uuids, int n, int out[3]) {
  try {
    std::vector<UUID> v(n);
    for (int i = 0; i < n; ++i) std::memcpy(v[i].data.data(), uuids + 16 * i, 16);
    const Transaction t = marginalizeVariables("MARGINALIZATION", v, g->graph);
    UpdateGraph(g->graph, t);
    out[0] = static_cast<int>(t.removed_constraints.size()); out[1] = static_cast<int>(t.added_constraints.size()); out[2] = static_cast<int>(t.removed_variables.size());
    return 0;
  } catch (const std::out_of_range& e) { g->error = e.what(); return -1; }
  catch (const std::invalid_argument& e) { g->error = e.what(); return -2; }
  catch (const std::exception& e) { g->error = e.what(); return -3; }
}
// detail::marginalizeNext on flat arrays (all blocks three columns wide): term t has term_rows[t] rows and term_nvars[t]
// variables; vars, A (rows x 3 per variable, row-major) and b are the terms' data one after the other. Outputs sized by
// the caller (at most 16 variables, 64 rows). Returns the number of rows of the marginal.
int bsfuse_marginalize_next(int nterms, const int* term_rows, const int* term_nvars, const int* vars, const double* A, const double* b,
                            int* out_nvars, int* out_vars, double* out_A, double* out_b) {
  std::vector<LinearTerm> terms(nterms);
  for (int t = 0; t < nterms; ++t) {
    for (int i = 0; i < term_nvars[t]; ++i) { terms[t].variables.push_back(static_cast<unsigned int>(*vars++)); terms[t].cols.push_back(3); terms[t].A.emplace_back(A, A + 3 * term_rows[t]); A += 3 * term_rows[t]; }
    terms[t].b.assign(b, b + term_rows[t]); b += term_rows[t];
  }
  const LinearTerm m = marginalizeNext(terms);
  *out_nvars = static_cast<int>(m.variables.size());
  for (size_t i = 0; i < m.variables.size(); ++i) { out_vars[i] = static_cast<int>(m.variables[i]); std::memcpy(out_A, m.A[i].data(), sizeof(double) * m.A[i].size()); out_A += m.A[i].size(); }
  std::memcpy(out_b, m.b.data(), sizeof(double) * m.b.size());
  return static_cast<int>(m.b.size());
}

// ---- sliding-window replay (smoother_replay.h) ----
bsfuse_replay* bsfuse_replay_new(int total_keyframes, int window_keyframes, uint64_t seed) {
  auto* h = new bsfuse_replay();
  try { h->r.reset(new SmootherReplay(total_keyframes, window_keyframes, seed)); }
  catch (const std::exception& e) { h->error = e.what(); }
  return h;
}
void bsfuse_replay_free(bsfuse_replay* h) { delete h; }
const char* bsfuse_replay_error(bsfuse_replay* h) { return h->error.c_str(); }
int bsfuse_replay_done(bsfuse_replay* h) { return !h->r || h->r->Done(); }
// steps 1-2 of a smoother cycle (transaction + pseudo-marginalisation); returns host seconds, < 0 on error
double bsfuse_replay_advance(bsfuse_replay* h) {
  try { return h->r->Advance(); } catch (const std::exception& e) { h->error = e.what(); return -1.0; }
}
const bsnls_problem* bsfuse_replay_flatten(bsfuse_replay* h) {
  try { return &h->r->graph().flattened().c; } catch (const std::exception& e) { h->error = e.what(); return nullptr; }
}
long long bsfuse_replay_observations(bsfuse_replay* h) { return h->r->num_observations(); }
int bsfuse_replay_optimize(bsfuse_replay* h, const bsnls_options* o, bsnls_summary* s, bsnls_iteration* its, int max_its, double timing[5]) {
  try {
    const SolverSummary r = h->r->graph().optimize(FromC(o));
    ToC(r, s, its, max_its);
    if (timing) TimingOut(h->r->graph(), timing);
    return 0;
  } catch (const std::invalid_argument& e) { h->error = e.what(); return -1; }
  catch (const std::exception& e) { h->error = e.what(); return -3; }
}
void bsfuse_replay_current_values(bsfuse_replay* h, double* out) {   // in the order of the last bsfuse_replay_flatten
  const FlatProblem& f = h->r->graph().lastProblem();
  for (size_t b = 0; b < f.block_variable.size(); ++b)
    std::memcpy(out + f.block_offset[b], f.block_variable[b]->data(), sizeof(double) * f.block_size[b]);
}
int bsfuse_evaluate(bsfuse_graph* g, double* cost) {
  try { return g->graph.evaluate(cost, nullptr, nullptr) ? 0 : -3; }
  catch (const std::exception& e) { g->error = e.what(); return -3; }
}

// bs_common::PreIntegrator (host/preintegrator.cpp) on a caller-supplied IMU stream: rows (t, w[3], a[3]) sorted by time, all
// with t <= t_end (ImuPreintegration keeps its buffer that way). bg / ba: bias estimates subtracted from the measurements
// and stored as the factor's linearisation point. noise = (cov_gyro, cov_accel, cov_gyro_bias, cov_accel_bias).
// delta_out: t, q (w,x,y,z), p, v; consts: the BSNLS_IMU_DELTA_STRIDE block of include/bsnls.h.
int bsfuse_preintegrate(const double* imu7, int n, double t_end, const double bg[3], const double ba[3], const double noise[4],
                        double info_weight, double delta_out[11], double consts[288]) {
  if (!imu7 || n <= 0) return -1;
  PreIntegrator pre; pre.Reset();
  pre.cov_w = noise[0]; pre.cov_a = noise[1]; pre.cov_bg = noise[2]; pre.cov_ba = noise[3];
  for (int i = 0; i < n; ++i) { IMUData d; d.t = imu7[7 * i]; for (int k = 0; k < 3; ++k) { d.w[k] = imu7[7 * i + 1 + k]; d.a[k] = imu7[7 * i + 4 + k]; } pre.data.push_back(d); }
  if (!pre.Integrate(t_end, bg, ba, true, true, true)) return -1;
  delta_out[0] = pre.delta.t; std::memcpy(delta_out + 1, pre.delta.q, 32); std::memcpy(delta_out + 5, pre.delta.p, 24); std::memcpy(delta_out + 8, pre.delta.v, 24);
  pre.FillFactorConstants(bg, ba, info_weight, consts);
  return 0;
}

// ---- host-logic probes for the CPU tests ----
void bsfuse_uuid_name(const unsigned char ns[16], const void* data, size_t n, unsigned char out[16]) {
  UUID u; std::memcpy(u.data.data(), ns, 16);
  const UUID r = uuid::name_uuid(u, data, n);
  std::memcpy(out, r.data.data(), 16);
}
void bsfuse_uuid_stamped(const char* type, uint32_t sec, uint32_t nsec, unsigned char out[16]) {
  const UUID r = uuid::generate(type, sec, nsec, uuid::nil());
  std::memcpy(out, r.data.data(), 16);
}
void bsfuse_uuid_landmark(const char* type, uint64_t id, unsigned char out[16]) {
  const UUID r = uuid::generate(type, id);
  std::memcpy(out, r.data.data(), 16);
}
uint64_t bsfuse_uuid_hash(const unsigned char in[16]) { UUID u; std::memcpy(u.data.data(), in, 16); return UuidHash()(u); }

// Container semantics of fuse_graphs::HashGraph (hash_graph.cpp:100-301): returns 0 when every check holds,
// else the number of the failed check.
int bsfuse_selftest_containers() {
  GpuHashGraph g;
  auto q = variables::Orientation3DStamped(Time(1, 0)); auto p = variables::Position3DStamped(Time(1, 0));
  auto lm = variables::Point3DLandmark(7);
  if (!g.addVariable(q) || !g.addVariable(p)) return 1;
  const double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, K[9] = {100, 0, 50, 0, 100, 50, 0, 0, 1}, px[2] = {50, 50};
  auto c = constraints::EuclideanReprojection("t", *q, *p, *lm, T, K, px, 1.0);
  try { g.addConstraint(c); return 2; } catch (const std::logic_error&) {}       // unknown variable
  g.addVariable(lm);
  if (!g.addConstraint(c)) return 3;
  if (g.addConstraint(c)) return 4;                                               // duplicate
  try { g.removeVariable(lm->uuid()); return 5; } catch (const std::logic_error&) {}  // still in use
  try { g.getVariable(variables::Point3DLandmark(8)->uuid()); return 6; } catch (const std::out_of_range&) {}
  try { g.getConstraint(NextConstraintUuid()); return 7; } catch (const std::out_of_range&) {}
  if (g.getConnectedConstraints(lm->uuid()).size() != 1) return 8;
  g.holdVariable(p->uuid());
  if (!g.isVariableOnHold(p->uuid())) return 9;
  FlatProblem f; Flatten(g, &f);
  int held = 0; for (auto cst : f.block_constant) held += cst;
  if (held != 1 || f.c.num_blocks != 3 || f.cgroups.size() != 1) return 10;
  if (!g.removeConstraint(c->uuid()) || !g.removeVariable(lm->uuid())) return 11;
  SolverOptions bad; bad.max_num_iterations = -1;
  try { g.optimize(bad); return 12; } catch (const std::invalid_argument&) {}
  auto unknown = std::make_shared<Constraint>(NextConstraintUuid(), "fuse_constraints::AbsolutePose2DStampedConstraint", "t",
                                              std::vector<UUID>{p->uuid()}, std::vector<double>{});
  g.addConstraint(unknown);
  try { Flatten(g, &f); return 13; } catch (const std::invalid_argument&) {}
  return 0;
}

}  // extern "C"
