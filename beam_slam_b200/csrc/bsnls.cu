// C-ABI implementation (include/bsnls.h): host-side structure analysis of the flattened graph, device buffers,
// and the enqueue sequence of the LM loop. The arithmetic lives in kernels.cuh; nothing here computes on the CPU.
#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: ranges around the C-ABI calls and the phases of a pass (nsys / ncu --nvtx)

#include "kernels.cuh"
#include "tile_chol.cuh"
#include "nccl_dyn.h"

using namespace bsnls;

namespace {

struct NvtxRange { explicit NvtxRange(const char* name) { nvtxRangePushA(name); } ~NvtxRange() { nvtxRangePop(); } };

thread_local std::string g_create_error;

const int kNumBlocks[BSNLS_NUM_FACTOR_TYPES] = {3, 10, 5, 4, 6, 2, 2, 1, 2, BSNLS_MARGINAL_MAX_BLOCKS, 1, 1};   // block slots per factor in a group's `blocks`
const int kNumRes[BSNLS_NUM_FACTOR_TYPES] = {2, 15, 15, 6, 6, 6, 3, 3, 3, BSNLS_MARGINAL_MAX_ROWS, 2, 3};
const int kStride[BSNLS_NUM_FACTOR_TYPES] = {BSNLS_REPROJECTION_STRIDE, BSNLS_IMU_DELTA_STRIDE, BSNLS_IMU_PRIOR_STRIDE,
                                             BSNLS_POSE_STRIDE, BSNLS_POSE_STRIDE, BSNLS_POSE_STRIDE,
                                             BSNLS_ORIENTATION_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_MARGINAL_STRIDE, BSNLS_GRAVITY_STRIDE, BSNLS_ORIENTATION_STRIDE};
// blocks factor i of a group really uses: the marginal prior lists its blocks first and pads the slots with -1
inline int UsedBlocks(const bsnls_factor_group& G, int64_t i) {
  const int nbk = kNumBlocks[G.type];
  if (G.type != BSNLS_FACTOR_MARGINAL) return nbk;
  int n = 0; while (n < nbk && G.blocks[i * nbk + n] >= 0) ++n;
  return n;
}

// Device memory of a handle. bsnls_set_problem carves every buffer of a problem out of ONE device arena and fills the
// uploaded ones from ONE pinned staging buffer with a single host-to-device copy (a sliding window re-flattens its graph
// every smoother cycle: ~100 cudaMalloc + ~70 pageable copies per cycle cost more than the solve). While an arena is
// recording (g_rec != nullptr) DevMem::alloc / upload only reserve offsets; Arena::Commit sizes the arena (growing it
// with 25 % head room, so later windows of similar size allocate nothing), resolves the pointers and issues the copy.
struct DevMem;
struct Arena {
  char* base = nullptr; size_t cap = 0;          // device
  char* stage = nullptr; size_t stage_cap = 0;   // pinned host mirror of the uploaded region [0, up_used)
  size_t up_used = 0, work_used = 0;
  std::vector<DevMem*> members;
  struct Deferred { DevMem* m; const void* src; size_t bytes; };
  std::vector<Deferred> deferred;   // large arrays skip the staging mirror: copied straight from the caller's vector at Commit
  static size_t Align(size_t n) { return (n + 255) & ~size_t(255); }
  void Begin() { up_used = 0; work_used = 0; members.clear(); deferred.clear(); }
  cudaError_t StageReserve(size_t n) {
    if (n <= stage_cap) return cudaSuccess;
    const size_t nc = std::max(std::max(2 * stage_cap, n + n / 4), size_t(1) << 20);
    char* ns = nullptr;
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&ns), nc, cudaHostAllocDefault);
    if (e != cudaSuccess) return e;
    if (stage) { std::memcpy(ns, stage, up_used); cudaFreeHost(stage); }
    stage = ns; stage_cap = nc;
    return cudaSuccess;
  }
  cudaError_t Commit(cudaStream_t s);
  void Release() { if (base) cudaFree(base); if (stage) cudaFreeHost(stage); base = nullptr; stage = nullptr; cap = stage_cap = 0; }
};
thread_local Arena* g_rec = nullptr;

struct DevMem {
  void* p = nullptr; size_t bytes = 0;
  bool view = false, in_work = false; size_t off = 0;   // view: inside the handle's arena (not owned)
  ~DevMem() { if (p && !view) cudaFree(p); }
  cudaError_t alloc(size_t n) {
    if (n == 0) n = 8;
    if (g_rec) {   // reserve a slot of the arena's work region
      if (p && !view) cudaFree(p);
      p = nullptr; view = true; in_work = true; bytes = n; off = g_rec->work_used; g_rec->work_used += Arena::Align(n);
      g_rec->members.push_back(this);
      return cudaSuccess;
    }
    if (p && !view && n <= bytes) return cudaSuccess;
    if (p && !view) cudaFree(p);
    p = nullptr; bytes = 0; view = false;
    cudaError_t e = cudaMalloc(&p, n);
    if (e == cudaSuccess) bytes = n;
    return e;
  }
  template <typename T> T* as() const { return static_cast<T*>(p); }
};
cudaError_t Arena::Commit(cudaStream_t s) {
  const size_t up = Align(up_used), need = up + work_used;
  if (need > cap) {
    if (base) { cudaError_t e0 = cudaStreamSynchronize(s); if (e0 != cudaSuccess) return e0; cudaFree(base); base = nullptr; cap = 0; }
    const size_t nc = need + need / 4 + (size_t(1) << 16);
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&base), nc);
    if (e != cudaSuccess) return e;
    cap = nc;
  }
  for (DevMem* m : members) m->p = base + (m->in_work ? up + m->off : m->off);
  members.clear();
  cudaError_t e = up_used ? cudaMemcpyAsync(base, stage, up_used, cudaMemcpyHostToDevice, s) : cudaSuccess;
  for (const Deferred& d : deferred) if (e == cudaSuccess) e = cudaMemcpyAsync(d.m->p, d.src, d.bytes, cudaMemcpyHostToDevice, s);   // pageable source: returns once staged
  deferred.clear();
  return e;
}

}  // namespace

struct bsnls_handle {
  bsnls_options opt{};
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  cudaStream_t side = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;   // generic factors run beside the reprojection pass
  cudaEvent_t pev[10] = {};   // per-kernel marks for bsnls_last_breakdown_ms
  bool profiling = false; float breakdown[8] = {};
  std::string error;
  int64_t launches = 0;
  bool has_problem = false, jac_valid = false;
  // one LM pass / the closing Jacobian pass captured as CUDA graphs (single GPU): the kernel sequence is static, every
  // kernel is gated by the device state, so replaying the graph replaces ~40 stream launches per iteration
  int64_t graph_pass_launches = 0, graph_tail_launches = 0; const void* graph_recs = nullptr;
  cudaGraphExec_t graph_pass = nullptr, graph_tail = nullptr; bool graphs_valid = false; bool use_graphs = true;
  bool nccl_warm = false;        // the communicator has run its first collectives (uncaptured)
  cudaGraphExec_t graph_rp = nullptr; bool graph_rp_valid = false; int64_t graph_rp_launches = 0;   // bsnls_run_pass as one graph
  bool breakdown_enabled = false;   // bsnls_enable_breakdown: direct launches with CUDA events between the kernels
  int solves_on_structure = 0;   // graphs are captured from the second solve of one structure on: capture + instantiation cost more than one solve's launches
  float last_eval_ms = 0, last_schur_ms = 0;
  int rank = 0, world = 1; NcclComm comm = nullptr;
  int ne_global = 0, nobs_global = 0, nb_act_e = 0; DevMem d_owned, d_flag;
  std::vector<int> shard_e;   // this rank's landmarks as positions in the elimination order (bsnls_get_shard)

  // host-side structure
  int num_blocks = 0; int64_t num_values = 0;
  std::vector<int> blk_off, blk_size, blk_tan, blk_manifold, active_blocks, eblocks, fblocks;
  int ne = 0, nf = 0, np = 0, nobs = 0, ngf = 0, ngf_active = 0, npairs = 0, ngdest = 0, G = 8;
  int64_t nitems = 0;
  std::vector<int> pose_q_blk, pose_t_blk;
  std::vector<int> pair_a, pair_b;
  std::vector<int> gd_row, gd_col;
  bool tiled = false; std::vector<int> f_api;   // tile-sparse Cholesky; internal f position -> API (program-order) position
  TileChol TC{}; std::vector<int> tc_col_ptr, tc_blk_slot; int tc_nT = 0; DevMem d_tc_blk_slot, d_tc_slot_blk, d_tc_row_ptr, d_tc_row_tile, d_tc_braw, d_tc_P, d_tc_tile_off, d_tc_col_ptr, d_tc_col_rows, d_tc_T, d_tc_W, d_tc_b; size_t tc_T_bytes = 0;
  struct TcLevel { int col_begin, ncols, pan_begin, npan, npan_tiles, upd_begin, nupd, fw_begin, nfw, bw_begin, nbw; };   // one level of the elimination tree
  std::vector<TcLevel> tc_levels;
  DevMem d_tc_pan_k, d_tc_pan_x, d_tc_upd_ptr, d_tc_upd_fac, d_tc_upd_dst, d_tc_upd_a, d_tc_upd_b, d_tc_fw_ptr, d_tc_fw_row, d_tc_fw_tail, d_tc_fw_src, d_tc_fw_k, d_tc_lev_cols, d_tc_bw_ptr, d_tc_bw_col, d_tc_bw_tail, d_tc_bw_src, d_tc_bw_i, d_tc_roots; int tc_nroots = 0;
  long long nlow = 0; DevMem d_low_slot, d_low_slot_t, d_xsend, d_xrecv;   // sharded solve: lower-triangular blocks of S, packed exchange buffers
  int chunk_max_slots = 1; DevMem d_syrk_plan, d_chunk_tile_off, d_pp_toff, d_pp_info;
  int n_local_pairs = 0; DevMem d_local_pairs;   // pairs with contributions from this rank's landmarks
  bool fused = false; DevMem d_chunk_pose, d_chunk_stage_ptr, d_stage_lm, d_pc_ptr, d_pc_idx;   // chunk-major fused pass
  int nchunks = 0; DevMem d_obs_slot, d_chunk_end, d_chunk_begin, d_chunk_nslots, d_pp_ptr, d_part;
  bool sparse = false, pcg = false; int64_t nnzb = 0; std::vector<int> bsr_rowptr, bsr_col; int pcg_grid = 0;
  // residual-block bookkeeping for bsnls_evaluate (program order)
  struct ResRef { int type; int idx; int row; };   // idx: sorted-observation id or generic-factor id
  std::vector<ResRef> res_order; int64_t num_res_rows = 0;
  int64_t gfJ_size = 0, gfR_size = 0;
  std::vector<int> gf_roff_h, gf_nres_h, gf_joff_h, gf_nblk_h;
  int nb_obs = 0, nb_lm = 0, nb_act = 0;  // partial-sum block counts
  int n_eval_items = 0, nb_eval = 0;      // pose-major Jacobian pass: warp items, blocks of 8 items (cost partials at x)
  std::vector<int> obs_pm_h; DevMem d_obs_pm;
  DevMem d_eval_items, d_eval_item_ptr, d_pm_lm_off, d_pm_cam, d_pm_pixel, d_pose_tab, d_pose_part;

  DeviceProblem P{};
  Arena arena;   // every buffer bsnls_set_problem creates lives here
  // device memory
  DevMem d_blk_off, d_blk_size, d_blk_tan, d_blk_manifold, d_active, d_pose_q_off, d_pose_t_off, d_pose_q_tan,
      d_pose_t_tan, d_pose_ptr, d_pose_obs, d_obs_pose, d_obs_lm, d_obs_cam, d_obs_pixel, d_lm_ptr, d_lm_off, d_cams,
      d_gf_type, d_gf_nblk, d_gf_nres, d_gf_blocks, d_gf_coff, d_gf_consts, d_gf_loss_type, d_gf_loss_a, d_gf_joff,
      d_gf_roff, d_pair_a, d_pair_b, d_pair_ptr, d_item_i, d_item_j, d_fg_ptr, d_fg_fac, d_fg_lb, d_gd_row, d_gd_col,
      d_gd_ptr, d_gd_fac, d_gd_bi, d_gd_bj, d_is_pose, d_bsr_rowptr, d_bsr_col, d_diag_slot, d_pair_slot, d_gd_slot,
      d_pcg_x, d_pcg_r, d_pcg_z, d_pcg_p, d_pcg_Ap, d_pcg_Minv, d_pcg_part, d_pcgd_zsend, d_pcgd_zfull, d_pcgd_xsend, d_pcgd_scal, d_pcgd_state;
  int pcgd_rb0 = 0, pcgd_rb1 = 0; int* pcgd_done_host = nullptr;   // sharded PCG: own block rows, pinned copy of the done flag
  DevMem d_lmz, d_gradL, d_colsqL, d_gradS, d_colsqS, d_zero;
  cudaEvent_t ev_stage = nullptr; bool stage_busy = false;
  DevMem d_x, d_xc, d_x0, d_Jrec, d_Vrec, d_minv, d_grad, d_colsq, d_scale, d_y, d_yf, d_rhs, d_S, d_poseH, d_gfJ,
      d_gfR, d_cost_part, d_cand_part, d_model_part, d_step_part, d_candsq_part, d_gsq_part, d_gmax_part, d_xsq_part,
      d_red, d_state, d_recs;
  int max_records = 0;

  int fail(int code, const char* fmt, ...) {
    char buf[512]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    error = buf; return code;
  }
};

#define CK(expr)                                                                                      \
  do { cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e)); } while (0)
#define MARK(h, i) do { if ((h)->profiling) cudaEventRecord((h)->pev[i], (h)->stream); } while (0)
#define LAUNCH_SMEM(h, kernel, grid, block, smem, ...)                                                 \
  do { kernel<<<(grid), (block), (smem), (h)->stream>>>(__VA_ARGS__); (h)->launches++; } while (0)
#define LAUNCH(h, kernel, grid, block, ...)                                                            \
  do { kernel<<<(grid), (block), 0, (h)->stream>>>(__VA_ARGS__); (h)->launches++; } while (0)

namespace {

constexpr size_t kDeferBytes = size_t(1) << 20;
template <typename T> cudaError_t upload(DevMem& m, const std::vector<T>& v, cudaStream_t s) {
  if (g_rec && v.size() * sizeof(T) >= kDeferBytes) {   // large: no second host copy; v must outlive Arena::Commit (function-scope vectors of set_problem)
    cudaError_t e = m.alloc(v.size() * sizeof(T));
    if (e == cudaSuccess) g_rec->deferred.push_back({&m, v.data(), v.size() * sizeof(T)});
    return e;
  }
  if (g_rec) {   // copy into the pinned staging mirror; the device pointer is resolved by Arena::Commit
    const size_t n = std::max<size_t>(v.size() * sizeof(T), 8);
    cudaError_t e = g_rec->StageReserve(g_rec->up_used + Arena::Align(n));
    if (e != cudaSuccess) return e;
    if (m.p && !m.view) cudaFree(m.p);
    m.p = nullptr; m.view = true; m.in_work = false; m.bytes = n; m.off = g_rec->up_used;
    if (!v.empty()) std::memcpy(g_rec->stage + m.off, v.data(), v.size() * sizeof(T));
    g_rec->up_used += Arena::Align(n);
    g_rec->members.push_back(&m);
    return cudaSuccess;
  }
  cudaError_t e = m.alloc(v.size() * sizeof(T));
  if (e != cudaSuccess || v.empty()) return e;
  return cudaMemcpyAsync(m.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s);
}

Options MakeOptions(const bsnls_options& o) {
  Options r;
  r.min_lm_diagonal = o.min_lm_diagonal; r.max_lm_diagonal = o.max_lm_diagonal;
  r.function_tolerance = o.function_tolerance; r.gradient_tolerance = o.gradient_tolerance;
  r.parameter_tolerance = o.parameter_tolerance; r.max_radius = o.max_trust_region_radius;
  r.min_radius = o.min_trust_region_radius; r.min_relative_decrease = o.min_relative_decrease;
  r.max_num_iterations = o.max_num_iterations; r.max_invalid = o.max_num_consecutive_invalid_steps;
  r.jacobi_scaling = o.jacobi_scaling;
  return r;
}

bool ValidOptions(const bsnls_options& o, std::string* why) {
  auto bad = [&](const char* m) { *why = m; return false; };
  if (o.max_num_iterations < 0) return bad("max_num_iterations >= 0 violated");
  if (o.max_solver_time_in_seconds < 0) return bad("max_solver_time_in_seconds >= 0 violated");
  if (o.function_tolerance < 0 || o.gradient_tolerance < 0 || o.parameter_tolerance < 0) return bad("tolerances >= 0 violated");
  if (!(o.initial_trust_region_radius > 0) || !(o.min_trust_region_radius > 0) || !(o.max_trust_region_radius > 0)) return bad("trust region radii > 0 violated");
  if (o.min_trust_region_radius > o.max_trust_region_radius || o.initial_trust_region_radius > o.max_trust_region_radius ||
      o.min_trust_region_radius > o.initial_trust_region_radius) return bad("trust region radius ordering violated");
  if (o.min_relative_decrease < 0) return bad("min_relative_decrease >= 0 violated");
  if (o.min_lm_diagonal < 0 || o.min_lm_diagonal > o.max_lm_diagonal) return bad("lm_diagonal bounds violated");
  if (o.max_num_consecutive_invalid_steps < 0) return bad("max_num_consecutive_invalid_steps >= 0 violated");
  if (o.linear_solver < 0 || o.linear_solver > BSNLS_LINEAR_AUTO) return bad("unknown linear_solver");
  return true;
}

const char* MessageText(int id) {
  switch (id) {
    case MSG_MAX_ITER: return "Maximum number of iterations reached.";
    case MSG_GRAD_TOL: return "Gradient tolerance reached.";
    case MSG_MIN_RADIUS: return "Minimum trust region radius reached.";
    case MSG_PARAM_TOL: return "Parameter tolerance reached.";
    case MSG_FUNC_TOL: return "Function tolerance reached.";
    case MSG_INVALID_STEPS: return "Number of consecutive invalid steps more than Solver::Options::max_num_consecutive_invalid_steps.";
    case MSG_MAX_TIME: return "Maximum solver time reached.";
    case MSG_NONFINITE: return "Residual and Jacobian evaluation failed (non-finite cost).";
    default: return "";
  }
}

int BlocksFor(int n, int per) { return std::max(1, (n + per - 1) / per); }

// Host-side structure analysis of a refinement-sized graph (millions of observations) runs in slices on host threads;
// sliding windows (below `grain` items per thread) stay on the calling thread. fn(begin, end) must only write its slice.
template <class F> void ParallelFor(int n, int grain, F fn) {
  int nt = (int)std::min<long long>(std::min(16u, std::max(1u, std::thread::hardware_concurrency())), (long long)n / std::max(1, grain));
  if (nt <= 1) { fn(0, n); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) th.emplace_back([&, t]() { fn((int)((long long)n * t / nt), (int)((long long)n * (t + 1) / nt)); });
  for (auto& x : th) x.join();
}

// ---- enqueue helpers (all asynchronous on h->stream) ----
// ---- NCCL exchange of the landmark-sharded solve (no-ops on a single GPU) ----
int AllReduce(bsnls_handle* h, double* buf, size_t count, int op) {
  if (h->world <= 1 || count == 0) return BSNLS_OK;
  const int rc = Nccl().AllReduce(buf, buf, count, kNcclFloat64, op, h->comm, h->stream);
  return rc == kNcclSuccess ? BSNLS_OK : h->fail(BSNLS_ERR_CUDA, "ncclAllReduce: %s", Nccl().GetErrorString(rc));
}
int AllReduceScalars(bsnls_handle* h) {  // rank-local partial sums and the local max of red[]
  if (h->world <= 1) return BSNLS_OK;
  Nccl().GroupStart();
  int rc = AllReduce(h, h->d_red.as<double>(), RED_NUM, kNcclSum);
  if (rc == BSNLS_OK) rc = AllReduce(h, h->d_red.as<double>() + kRedLocalMax, 1, kNcclMax);
  Nccl().GroupEnd();
  return rc;
}
RedDesc Desc(const double* src, int n, int op, int slot, bool replicated) { return RedDesc{src, n, op, op ? (replicated ? kRedReplMax : kRedLocalMax) : slot + (replicated ? RED_NUM : 0)}; }

// chunk-major fused pass: Jacobians (or their stored records), landmark elimination and, with_system, the chunk tiles
void EnqueueFused(bsnls_handle* h, const Options& o, int with_system, int force_load) {
  LAUNCH_SMEM(h, chunk_fused_kernel, h->nchunks, 256, fused_smem_bytes(h->chunk_max_slots), h->P, h->d_x.as<double>(), h->d_pose_tab.as<double>(), h->d_Jrec.as<double>(),
              h->d_minv.as<double>(), h->d_grad.as<double>(), h->d_colsq.as<double>(), h->d_scale.as<double>(), h->d_lmz.as<double>(), h->d_part.as<double>(),
              h->d_pose_part.as<double>(), h->d_cost_part.as<double>(), o, h->d_state.as<LMState>(), with_system, force_load);
}
// phase A: residuals + Jacobians + cost at x. fused_system: in the fused mode, also eliminate and accumulate the tiles
void EnqueueEval(bsnls_handle* h, int fused_system = 0) {
  NvtxRange nvtx_range("enqueue: residuals + Jacobians");
  const DeviceProblem& P = h->P;
  LMState* st = h->d_state.as<LMState>();
  MARK(h, 0);
  if (h->fused) {
    const Options o = MakeOptions(h->opt);
    const bool fork = h->ngf > 0;
    if (fork) {
      cudaEventRecord(h->ev_fork, h->stream); cudaStreamWaitEvent(h->side, h->ev_fork, 0);
      eval_generic_kernel<0><<<BlocksFor(h->ngf, 4), 128, 0, h->side>>>(P, h->d_x.as<double>(), h->d_gfJ.as<double>(), h->d_gfR.as<double>(), h->d_cost_part.as<double>() + h->nb_eval, st);
      h->launches++;
      cudaEventRecord(h->ev_join, h->side);
    }
    LAUNCH(h, pose_prep_kernel, BlocksFor(h->np, 128), 128, P, h->d_x.as<double>(), h->d_pose_tab.as<double>(), st);
    EnqueueFused(h, o, fused_system, 0);
    MARK(h, 1);
    if (fork) cudaStreamWaitEvent(h->stream, h->ev_join, 0);
    MARK(h, 2);
    return;
  }
  // the generic factors (IMU, priors, relative poses) are independent of the reprojection pass until the gradient is
  // put together: they run on a side stream (fork / join by events, which also works under stream capture)
  const bool fork = h->ngf && h->nobs;
  if (fork) {
    cudaEventRecord(h->ev_fork, h->stream); cudaStreamWaitEvent(h->side, h->ev_fork, 0);
    eval_generic_kernel<0><<<BlocksFor(h->ngf, 4), 128, 0, h->side>>>(P, h->d_x.as<double>(), h->d_gfJ.as<double>(), h->d_gfR.as<double>(), h->d_cost_part.as<double>() + h->nb_eval, st);
    h->launches++;
    cudaEventRecord(h->ev_join, h->side);
  }
  if (h->nobs) {
    LAUNCH(h, pose_prep_kernel, BlocksFor(h->np, 128), 128, P, h->d_x.as<double>(), h->d_pose_tab.as<double>(), st);
    LAUNCH(h, eval_reproj_pose_kernel, h->nb_eval, 256, P, h->d_x.as<double>(), h->d_pose_tab.as<double>(), h->d_Jrec.as<double>(), h->d_cost_part.as<double>(), h->d_pose_part.as<double>(), st);
  }
  MARK(h, 1);
  if (fork) cudaStreamWaitEvent(h->stream, h->ev_join, 0);
  else if (h->ngf) LAUNCH(h, eval_generic_kernel<0>, BlocksFor(h->ngf, 4), 128, P, h->d_x.as<double>(), h->d_gfJ.as<double>(), h->d_gfR.as<double>(), h->d_cost_part.as<double>() + h->nb_eval, st);
  MARK(h, 2);
}
template <int G> void LaunchLandmark(bsnls_handle* h, const Options& o) {
  LAUNCH(h, landmark_kernel<G>, BlocksFor(h->ne * G, 256), 256, h->P, h->d_Jrec.as<double>(), h->d_Vrec.as<double>(), h->d_minv.as<double>(),
         h->d_grad.as<double>(), h->d_colsq.as<double>(), h->d_scale.as<double>(), h->d_lmz.as<double>(), o, h->d_state.as<LMState>());
}
void EnqueueLandmark(bsnls_handle* h, const Options& o) {  // phase B
  if (!h->ne) return;
  if (h->fused) { EnqueueFused(h, o, 1, 1); return; }   // stored Jacobian records -> elimination + chunk tiles
  switch (h->G) { case 4: LaunchLandmark<4>(h, o); break; case 8: LaunchLandmark<8>(h, o); break;
                  case 16: LaunchLandmark<16>(h, o); break; default: LaunchLandmark<32>(h, o); }
}
// Phases C and D1 of a pass, ordered so that a sharded solve needs ONE exchange: everything that only needs this
// rank's observations first (pose gradient / column norms, cost and landmark-gradient norms, the shard's reduced
// system), then one grouped all-reduce, then the kernels that need the sums (generic-factor gradient, Jacobi scale,
// f-block gradient norms, generic rows of S, damping).
void EnqueueGradientLocal(bsnls_handle* h) {
  const DeviceProblem& P = h->P; LMState* st = h->d_state.as<LMState>();
  if (h->np) LAUNCH(h, pose_reduce_kernel, BlocksFor(h->np * 27, 128), 128, P, h->d_pose_part.as<double>(), h->fused ? h->d_pc_ptr.as<int>() : h->d_eval_item_ptr.as<int>(),
                    h->fused ? h->d_pc_idx.as<int>() : (const int*)nullptr, h->d_poseH.as<double>(), h->d_gradL.as<double>(), h->d_colsqL.as<double>(), st);
  if (h->world == 1) return;   // one GPU: a single norm / reduction launch after the generic-factor gradient does it all
  const int nbe = h->nb_act_e;
  if (nbe) LAUNCH(h, grad_norm_kernel, nbe, 256, P, h->d_x.as<double>(), h->d_grad.as<double>(), h->d_gsq_part.as<double>(), h->d_gmax_part.as<double>(), 0, st);
  RedList L{}; L.gate = 2; L.count = 0; L.keep = 0;
  if (h->nobs) L.d[L.count++] = Desc(h->d_cost_part.as<double>(), h->nb_eval, 0, RED_COST, false);
  if (nbe) L.d[L.count++] = Desc(h->d_gsq_part.as<double>(), nbe, 0, RED_GRAD_SQ, false);
  if (nbe) L.d[L.count++] = Desc(h->d_gmax_part.as<double>(), nbe, 1, 0, false);
  LAUNCH(h, reduce_kernel, 1, 256, L, h->d_red.as<double>(), st);
}
// with_system: the exchange of a full pass (reduced system + gradient + scalars); otherwise gradient + scalars only
int EnqueueExchange(bsnls_handle* h, bool with_system) {
  if (h->world <= 1) return BSNLS_OK;
  const DeviceProblem& P = h->P;
  // Pose gradient / column norms go out of place: the kernels that rebuild them are gated by the device state, the
  // collective is not, so it must be idempotent when they did not run. S and rhs are rebuilt on every pass; red[] is
  // only read by the (equally gated) controller.
  if (h->sparse) {   // one packed buffer, one collective (lower-triangular blocks of S only)
    const ExchangeLayout X{h->nlow * 9, P.nf_dim, h->world, h->rank, RED_NUM, kRedLocalMax, h->d_low_slot.as<int>(), h->d_low_slot_t.as<int>()};
    const long long total = X.nS + 3LL * P.nf_dim + RED_NUM + h->world + 1, first = with_system ? 0 : X.nS + P.nf_dim;
    const unsigned grid = (unsigned)((total - first + 255) / 256);
    LAUNCH(h, pack_exchange_kernel, grid, 256, X, first, h->d_S.as<double>(), h->d_rhs.as<double>(), h->d_gradL.as<double>(), h->d_colsqL.as<double>(), h->d_red.as<double>(), h->d_state.as<LMState>(), h->d_xsend.as<double>());
    const int rc1 = Nccl().AllReduce(h->d_xsend.as<double>() + first, h->d_xrecv.as<double>() + first, (size_t)(total - first), kNcclFloat64, kNcclSum, h->comm, h->stream);
    if (rc1 != kNcclSuccess) return h->fail(BSNLS_ERR_CUDA, "ncclAllReduce(exchange): %s", Nccl().GetErrorString(rc1));
    LAUNCH(h, unpack_exchange_kernel, grid, 256, X, first, h->d_xrecv.as<double>(), h->d_S.as<double>(), h->d_rhs.as<double>(), h->d_gradS.as<double>(), h->d_colsqS.as<double>(), h->d_red.as<double>(), h->d_state.as<LMState>());
    return BSNLS_OK;
  }
  Nccl().GroupStart();
  int rc = kNcclSuccess;
  if (with_system) {
    const size_t s_count = h->sparse ? (size_t)h->nnzb * 9 : (size_t)P.nf_dim * P.nf_dim;
    rc = Nccl().AllReduce(h->d_S.p, h->d_S.p, s_count, kNcclFloat64, kNcclSum, h->comm, h->stream);
    if (rc == kNcclSuccess) rc = Nccl().AllReduce(h->d_rhs.p, h->d_rhs.p, P.nf_dim, kNcclFloat64, kNcclSum, h->comm, h->stream);
  }
  if (rc == kNcclSuccess) rc = Nccl().AllReduce(h->d_gradL.p, h->d_gradS.p, P.nf_dim, kNcclFloat64, kNcclSum, h->comm, h->stream);
  if (rc == kNcclSuccess) rc = Nccl().AllReduce(h->d_colsqL.p, h->d_colsqS.p, P.nf_dim, kNcclFloat64, kNcclSum, h->comm, h->stream);
  if (rc == kNcclSuccess) rc = Nccl().AllReduce(h->d_red.p, h->d_red.p, RED_NUM, kNcclFloat64, kNcclSum, h->comm, h->stream);
  if (rc == kNcclSuccess) rc = Nccl().AllReduce(h->d_red.as<double>() + kRedLocalMax, h->d_red.as<double>() + kRedLocalMax, 1, kNcclFloat64, kNcclMax, h->comm, h->stream);
  Nccl().GroupEnd();
  return rc == kNcclSuccess ? BSNLS_OK : h->fail(BSNLS_ERR_CUDA, "ncclAllReduce(exchange): %s", Nccl().GetErrorString(rc));
}
void EnqueueGradientGlobal(bsnls_handle* h, const Options& o) {
  const DeviceProblem& P = h->P; LMState* st = h->d_state.as<LMState>();
  const double* gS = h->world > 1 ? h->d_gradS.as<double>() : h->d_gradL.as<double>();
  const double* cS = h->world > 1 ? h->d_colsqS.as<double>() : h->d_colsqL.as<double>();
  LAUNCH(h, gf_grad_kernel, BlocksFor(3 * h->nf, 128), 128, P, h->d_gfJ.as<double>(), h->d_gfR.as<double>(), gS, cS, h->d_is_pose.as<unsigned char>(),
         h->d_grad.as<double>(), h->d_colsq.as<double>(), h->d_scale.as<double>(), o, st);
  const int nbe = h->nb_act_e, nbf = h->nb_act - h->nb_act_e;
  const bool one = h->world == 1;
  if (one) LAUNCH(h, grad_norm_kernel, h->nb_act, 256, P, h->d_x.as<double>(), h->d_grad.as<double>(), h->d_gsq_part.as<double>(), h->d_gmax_part.as<double>(), 0, st);
  else if (nbf) LAUNCH(h, grad_norm_kernel, nbf, 256, P, h->d_x.as<double>(), h->d_grad.as<double>(), h->d_gsq_part.as<double>(), h->d_gmax_part.as<double>(), nbe, st);
  RedList L{}; L.gate = 2; L.count = 0; L.keep = one ? 0 : 1;   // sharded: replicated slots, added to what the exchange left in red[]
  if (one && h->nobs) L.d[L.count++] = Desc(h->d_cost_part.as<double>(), h->nb_eval, 0, RED_COST, false);
  if (one && nbe) L.d[L.count++] = Desc(h->d_gsq_part.as<double>(), nbe, 0, RED_GRAD_SQ, false);
  if (one && nbe) L.d[L.count++] = Desc(h->d_gmax_part.as<double>(), nbe, 1, 0, false);
  if (h->ngf_active) L.d[L.count++] = Desc(h->d_cost_part.as<double>() + h->nb_eval, h->ngf_active, 0, RED_COST, true);
  if (h->ngf > h->ngf_active) L.d[L.count++] = Desc(h->d_cost_part.as<double>() + h->nb_eval + h->ngf_active, h->ngf - h->ngf_active, 0, RED_FIXED, true);
  if (nbf) L.d[L.count++] = Desc(h->d_gsq_part.as<double>() + nbe, nbf, 0, RED_GRAD_SQ, true);
  if (nbf) L.d[L.count++] = Desc(h->d_gmax_part.as<double>() + nbe, nbf, 1, 0, true);
  LAUNCH(h, reduce_kernel, 1, 256, L, h->d_red.as<double>(), st);
}
void EnqueueAssembleLocal(bsnls_handle* h) {  // phase D1, this rank's share: S and rhs from its landmarks
  const DeviceProblem& P = h->P; LMState* st = h->d_state.as<LMState>();
  const size_t s_count = h->sparse ? (size_t)h->nnzb * 9 : (size_t)P.nf_dim * P.nf_dim;
  MARK(h, 5);
  cudaMemsetAsync(h->d_S.p, 0, sizeof(double) * s_count, h->stream);
  if (h->world > 1) cudaMemsetAsync(h->d_rhs.p, 0, sizeof(double) * P.nf_dim, h->stream);
  MARK(h, 6);
  if (h->nchunks && !h->fused) {
    LAUNCH_SMEM(h, chunk_syrk_kernel, h->nchunks, 256, syrk_smem_bytes(h->chunk_max_slots), P, h->d_Vrec.as<double>(), h->d_lmz.as<double>(), h->d_part.as<double>(), st);
  }
  // sharded solve: only the pairs this rank's landmarks touch (the rest of S keeps the zeros of the clear)
  const int* plist = h->world > 1 ? h->d_local_pairs.as<int>() : nullptr; const int npl = h->world > 1 ? h->n_local_pairs : h->npairs;
  if (npl && h->nchunks) LAUNCH(h, pair_reduce_kernel, BlocksFor(npl, 4), 144, P, h->d_Vrec.as<double>(), h->d_part.as<double>(), h->d_lmz.as<double>(), h->d_poseH.as<double>(), h->d_S.as<double>(), h->d_rhs.as<double>(), plist, npl, st);
  else if (npl) LAUNCH(h, pair_kernel, BlocksFor(npl, 4), 128, P, h->d_Vrec.as<double>(), h->d_lmz.as<double>(), h->d_poseH.as<double>(), h->d_grad.as<double>(), h->d_S.as<double>(), h->d_rhs.as<double>(), plist, npl, st);
  // sharded solve: the rows of the generic factors go into rank 0's share BEFORE the exchange (the sum over ranks then
  // holds them once), which takes their kernel off the critical path behind the collective
  if (h->world > 1 && h->rank == 0 && h->ngdest) LAUNCH(h, gf_schur_kernel, BlocksFor(h->ngdest, 16), 144, P, h->d_gfJ.as<double>(), h->d_S.as<double>(), st);
  MARK(h, 7);
}
void EnqueueAssembleFinish(bsnls_handle* h, const Options& o) {  // generic-factor rows of S (single GPU), gradient into rhs, damping
  const DeviceProblem& P = h->P; LMState* st = h->d_state.as<LMState>();
  if (h->ngdest && h->world == 1) LAUNCH(h, gf_schur_kernel, BlocksFor(h->ngdest, 16), 144, P, h->d_gfJ.as<double>(), h->d_S.as<double>(), st);
  LAUNCH(h, finish_system_kernel, BlocksFor(P.nf_dim, 256), 256, P, h->d_grad.as<double>(), h->d_colsq.as<double>(), h->d_scale.as<double>(),
         h->d_is_pose.as<unsigned char>(), h->d_S.as<double>(), h->d_rhs.as<double>(), o, st);
  MARK(h, 8);
}
// The Jacobian half of a pass (after EnqueueEval / EnqueueLandmark) up to the controller's input; with_system also
// assembles the reduced system. Kernel order is the same on one GPU, where the exchange is a no-op.
int EnqueueLinearise(bsnls_handle* h, const Options& o, bool with_system) {
  NvtxRange nvtx_range("enqueue: gradient + reduced-system assembly");
  EnqueueGradientLocal(h);
  if (h->world == 1 && with_system) {
    // one GPU: nothing to exchange, so the gradient / norm kernels (small, latency-bound) run on the side stream
    // while the main stream assembles the reduced system
    cudaEventRecord(h->ev_fork, h->stream); cudaStreamWaitEvent(h->side, h->ev_fork, 0);
    std::swap(h->stream, h->side);
    EnqueueGradientGlobal(h, o);
    std::swap(h->stream, h->side);
    cudaEventRecord(h->ev_join, h->side);
    EnqueueAssembleLocal(h);
    MARK(h, 9);
    cudaStreamWaitEvent(h->stream, h->ev_join, 0);
    return BSNLS_OK;
  }
  if (with_system) EnqueueAssembleLocal(h);
  const int rc = EnqueueExchange(h, with_system);
  if (rc != BSNLS_OK) return rc;
  MARK(h, 9);
  EnqueueGradientGlobal(h, o);
  return BSNLS_OK;
}
// factor = false: the tile factor of the previous call is reused (another right-hand side of the same matrix)
void EnqueueDenseSolve(bsnls_handle* h, bool factor = true) {  // phase D2: Cholesky + triangular solves
  NvtxRange nvtx_range("enqueue: Cholesky of the reduced system");
  const int n = h->P.nf_dim; LMState* st = h->d_state.as<LMState>();
  double* Sd = h->d_S.as<double>();
  if (h->tiled) {
    const TileChol& TC = h->TC;
    const size_t smem = sizeof(double) * kPanelSmemDoubles;
    if (factor) {
      cudaMemsetAsync(TC.T, 0, h->tc_T_bytes, h->stream);
      LAUNCH(h, bsr_to_tiles_kernel, (unsigned)((h->nnzb * 9 + (long long)TC.nT * kTS + 255) / 256), 256, h->P, TC, h->d_S.as<double>(), st);
      // the forward substitution is part of the panel kernels: the right-hand side goes into the tile slots first
      LAUNCH(h, tile_rhs_kernel, (TC.nT * kTS + 255) / 256, 256, TC, h->d_rhs.as<double>(), TC.braw, st);
      bool first_level = true;
      for (const auto& L : h->tc_levels) {   // factorisation, one level of the elimination tree at a time
        LAUNCH_SMEM(h, tile_panel_kernel, L.npan, 256, smem, TC, L.pan_begin, first_level ? 1 : 0, st);
        first_level = false;
        if (L.nupd > 0) LAUNCH_SMEM(h, tile_syrk_kernel, L.nupd, 256, smem, TC, L.upd_begin, st);
      }
    }
    // substitutions, level by level: forward (bottom-up; only for another right-hand side of a kept factor), backward (top-down)
    if (!factor) {
      LAUNCH(h, tile_rhs_kernel, (TC.nT * kTS + 255) / 256, 256, TC, h->d_rhs.as<double>(), TC.b, st);
      LAUNCH(h, tile_fwd_first_kernel, h->tc_levels[0].ncols, 256, TC, h->tc_levels[0].col_begin, st);
      for (const auto& L : h->tc_levels) if (L.nfw > 0) LAUNCH(h, tile_fwd_kernel, L.nfw, 256, TC, L.fw_begin, st);
    }
    LAUNCH(h, tile_bwd_first_kernel, h->tc_nroots, 256, TC, st);
    for (size_t lv = h->tc_levels.size(); lv-- > 0;) if (h->tc_levels[lv].nbw > 0) LAUNCH(h, tile_bwd_kernel, h->tc_levels[lv].nbw, 256, TC, h->tc_levels[lv].bw_begin, st);
    LAUNCH(h, tile_out_kernel, (TC.n + 255) / 256, 256, TC, h->d_yf.as<double>(), st);
    return;
  }
  // sliding windows (n <= kRbMaxN): the whole matrix in registers, one CTA, factorisation and both substitutions
  LAUNCH(h, chol_reg_kernel, 1, kRbThreads, Sd, n, h->d_rhs.as<double>(), h->d_yf.as<double>(), st);
}
// K7 in the sharded solve: block rows of S partitioned over the ranks (kernels.cuh, pcgd_*). Not enqueue-only: the host
// looks at the done flag every kPcgdPoll iterations (the flag is computed from all-reduced numbers, so every rank stops
// at the same iteration and the collectives stay matched).
int RunPartitionedPcg(bsnls_handle* h) {
  constexpr int kPcgdPoll = 8;
  const int n = h->P.nf_dim, rows = h->pcgd_rb1 - h->pcgd_rb0;
  const int nblk_e = std::max(1, BlocksFor(std::max(rows, 1), 256)), nblk_w = std::max(1, BlocksFor(std::max(rows, 1), 8));
  LMState* st = h->d_state.as<LMState>();
  PcgDistBuf B{h->d_pcgd_xsend.as<double>(), h->d_pcg_r.as<double>(), h->d_pcg_p.as<double>(), h->d_pcg_Ap.as<double>(), h->d_pcg_Minv.as<double>(),
               h->d_pcgd_zsend.as<double>(), h->d_pcgd_zfull.as<double>(), h->d_pcg_part.as<double>(), h->d_pcgd_scal.as<double>(),
               h->d_pcgd_state.as<PcgDistState>(), h->pcgd_rb0, h->pcgd_rb1, nblk_e, h->opt.pcg_max_iterations, h->opt.pcg_tolerance};
  const double* S = h->d_S.as<double>(); const double* rhs = h->d_rhs.as<double>();
  CK(cudaMemsetAsync(h->d_pcgd_zsend.p, 0, sizeof(double) * (n + 3), h->stream));
  CK(cudaMemsetAsync(h->d_pcgd_xsend.p, 0, sizeof(double) * n, h->stream));
  LAUNCH(h, pcgd_init_kernel, nblk_e, 256, h->P, S, rhs, B, st);
  LAUNCH(h, pcgd_sum_kernel, 1, 32, B.part, nblk_e, 3, B.zsend + n);
  int rc = Nccl().AllReduce(B.zsend, B.zfull, (size_t)n + 3, kNcclFloat64, kNcclSum, h->comm, h->stream);
  if (rc != kNcclSuccess) return h->fail(BSNLS_ERR_CUDA, "ncclAllReduce: %s", Nccl().GetErrorString(rc));
  LAUNCH(h, pcgd_start_kernel, BlocksFor(n, 256), 256, n, B);
  LAUNCH(h, pcgd_start_ctrl_kernel, 1, 1, n, B);
  B.nblk = nblk_w;
  PcgDistBuf Bu = B; Bu.nblk = nblk_e;
  for (int it = 1; it <= h->opt.pcg_max_iterations; ++it) {
    LAUNCH(h, pcgd_spmv_kernel, nblk_w, 256, h->P, S, B);
    LAUNCH(h, pcgd_sum_kernel, 1, 32, B.part, nblk_w, 1, B.scal);
    rc = Nccl().AllReduce(B.scal, B.scal, 1, kNcclFloat64, kNcclSum, h->comm, h->stream);
    if (rc != kNcclSuccess) return h->fail(BSNLS_ERR_CUDA, "ncclAllReduce: %s", Nccl().GetErrorString(rc));
    LAUNCH(h, pcgd_update_kernel, nblk_e, 256, Bu);
    LAUNCH(h, pcgd_sum_kernel, 1, 32, B.part, nblk_e, 2, B.zsend + n);
    rc = Nccl().AllReduce(B.zsend, B.zfull, (size_t)n + 2, kNcclFloat64, kNcclSum, h->comm, h->stream);
    if (rc != kNcclSuccess) return h->fail(BSNLS_ERR_CUDA, "ncclAllReduce: %s", Nccl().GetErrorString(rc));
    LAUNCH(h, pcgd_dir_kernel, BlocksFor(n, 256), 256, n, B);
    LAUNCH(h, pcgd_ctrl_kernel, 1, 1, n, B);
    if (it % kPcgdPoll == 0 || it == h->opt.pcg_max_iterations) {
      CK(cudaMemcpyAsync(h->pcgd_done_host, &h->d_pcgd_state.as<PcgDistState>()->done, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
      CK(cudaStreamSynchronize(h->stream));
      if (*h->pcgd_done_host) break;
    }
  }
  // every rank needs the whole reduced step for its landmarks' back-substitution
  rc = Nccl().AllReduce(B.x, h->d_yf.as<double>(), (size_t)n, kNcclFloat64, kNcclSum, h->comm, h->stream);
  if (rc != kNcclSuccess) return h->fail(BSNLS_ERR_CUDA, "ncclAllReduce: %s", Nccl().GetErrorString(rc));
  LAUNCH(h, pcgd_finish_kernel, 1, 1, B, st);
  return BSNLS_OK;
}
int EnqueueSparseSolve(bsnls_handle* h) {  // phase D2 (K7): cooperative PCG, result in yf
  if (h->world > 1 && std::getenv("BSNLS_PCG_REPLICATED") == nullptr) return RunPartitionedPcg(h);
  PcgBuffers B{h->d_yf.as<double>(), h->d_pcg_r.as<double>(), h->d_pcg_z.as<double>(), h->d_pcg_p.as<double>(), h->d_pcg_Ap.as<double>(),
               h->d_pcg_Minv.as<double>(), h->d_pcg_part.as<double>(), h->opt.pcg_max_iterations, h->opt.pcg_tolerance};
  DeviceProblem P = h->P; const double* S = h->d_S.as<double>(); const double* rhs = h->d_rhs.as<double>(); LMState* st = h->d_state.as<LMState>();
  void* args[] = {&P, &S, &rhs, &B, &st};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)pcg_kernel, dim3(h->pcg_grid), dim3(256), args, 0, h->stream);
  h->launches++;
  return e == cudaSuccess ? BSNLS_OK : h->fail(BSNLS_ERR_CUDA, "cooperative launch of pcg_kernel: %s", cudaGetErrorString(e));
}
template <int G> void LaunchBacksub(bsnls_handle* h) {
  LAUNCH(h, backsub_kernel<G>, h->nb_lm, 256, h->P, h->d_Jrec.as<double>(), h->d_minv.as<double>(), h->d_grad.as<double>(), h->d_yf.as<double>(),
         h->d_y.as<double>(), h->d_model_part.as<double>(), h->d_state.as<LMState>());
}
int EnqueueStep(bsnls_handle* h) {  // phase D3: back-substitution, model cost, candidate, candidate cost
  NvtxRange nvtx_range("enqueue: back-substitution + candidate");
  const DeviceProblem& P = h->P; LMState* st = h->d_state.as<LMState>();
  if (h->ne) switch (h->G) { case 4: LaunchBacksub<4>(h); break; case 8: LaunchBacksub<8>(h); break;
                             case 16: LaunchBacksub<16>(h); break; default: LaunchBacksub<32>(h); }
  LAUNCH(h, gf_model_kernel, BlocksFor(std::max(32 * h->ngf, std::min(P.nf_dim, 1 << 16)), 128), 128, P, h->d_gfJ.as<double>(), h->d_gfR.as<double>(), h->d_yf.as<double>(), h->d_y.as<double>(),
         h->d_model_part.as<double>() + h->nb_lm, st);
  LAUNCH(h, candidate_kernel, h->nb_act, 256, P, h->d_x.as<double>(), h->d_y.as<double>(), h->d_xc.as<double>(), h->d_step_part.as<double>(), h->d_candsq_part.as<double>(), st);
  if (h->nobs) LAUNCH(h, eval_reproj_cost_kernel, h->nb_obs, 256, P, h->d_xc.as<double>(), h->d_cand_part.as<double>(), st);
  if (h->ngf) LAUNCH(h, eval_generic_kernel<1>, BlocksFor(h->ngf, 4), 128, P, h->d_xc.as<double>(), nullptr, nullptr, h->d_cand_part.as<double>() + h->nb_obs, st);
  RedList L{}; L.gate = 1; L.count = 0;
  const int nbe = h->nb_act_e, nbf = h->nb_act - h->nb_act_e;
  if (h->nobs) L.d[L.count++] = Desc(h->d_cand_part.as<double>(), h->nb_obs, 0, RED_CAND_COST, false);
  if (h->ngf_active) L.d[L.count++] = Desc(h->d_cand_part.as<double>() + h->nb_obs, h->ngf_active, 0, RED_CAND_COST, true);
  if (h->ne) L.d[L.count++] = Desc(h->d_model_part.as<double>(), h->nb_lm, 0, RED_MODEL, false);
  if (h->ngf_active) L.d[L.count++] = Desc(h->d_model_part.as<double>() + h->nb_lm, h->ngf_active, 0, RED_MODEL, true);
  if (nbe) L.d[L.count++] = Desc(h->d_step_part.as<double>(), nbe, 0, RED_STEP_SQ, false);
  if (nbf) L.d[L.count++] = Desc(h->d_step_part.as<double>() + nbe, nbf, 0, RED_STEP_SQ, true);
  if (nbe) L.d[L.count++] = Desc(h->d_candsq_part.as<double>(), nbe, 0, RED_CAND_SQ, false);
  if (nbf) L.d[L.count++] = Desc(h->d_candsq_part.as<double>() + nbe, nbf, 0, RED_CAND_SQ, true);
  LAUNCH(h, reduce_kernel, 1, 256, L, h->d_red.as<double>(), st);
  return AllReduceScalars(h);
}
int EnqueueXNorm(bsnls_handle* h) {  // |x| over the reduced program's blocks -> st->x_norm
  LMState* st = h->d_state.as<LMState>();
  LAUNCH(h, xnorm_kernel, h->nb_act, 256, h->P, h->d_x.as<double>(), h->d_xsq_part.as<double>());
  RedList L{}; L.gate = 0; L.count = 0;
  const int nbe = h->nb_act_e, nbf = h->nb_act - h->nb_act_e;
  if (nbe) L.d[L.count++] = Desc(h->d_xsq_part.as<double>(), nbe, 0, RED_X_SQ, false);
  if (nbf) L.d[L.count++] = Desc(h->d_xsq_part.as<double>() + nbe, nbf, 0, RED_X_SQ, true);
  LAUNCH(h, reduce_kernel, 1, 256, L, h->d_red.as<double>(), st);
  const int rc = AllReduceScalars(h);
  LAUNCH(h, init_xnorm_kernel, 1, 1, st, h->d_red.as<double>());
  return rc;
}
// after a sharded solve every rank holds the complete parameter vector: owners keep their entries, others zero, sum
int GatherValues(bsnls_handle* h) {
  if (h->world <= 1) return BSNLS_OK;
  LAUNCH(h, mask_owned_kernel, BlocksFor((int)h->num_values, 256), 256, h->d_x.as<double>(), h->d_owned.as<unsigned char>(), (int)h->num_values);
  return AllReduce(h, h->d_x.as<double>(), (size_t)h->num_values, kNcclSum);
}

int InitState(bsnls_handle* h, bool keep_scale, double radius = 0.0) {
  LMState s{};
  s.radius = radius > 0.0 ? radius : h->opt.initial_trust_region_radius; s.decrease_factor = 2.0;
  s.need_jacobian = 1; s.scale_ready = keep_scale ? 1 : 0; s.max_records = h->max_records;
  cudaError_t e = cudaMemcpyAsync(h->d_state.p, &s, sizeof(s), cudaMemcpyHostToDevice, h->stream);
  if (e != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "state upload: %s", cudaGetErrorString(e));
  e = cudaMemsetAsync(h->d_red.p, 0, sizeof(double) * 32, h->stream);
  if (e != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "red memset: %s", cudaGetErrorString(e));
  return BSNLS_OK;
}

int EnqueueTail(bsnls_handle* h, const Options& o, bsnls_iteration* recs) {   // Jacobian pass that closes an accepted step (or IterationZero)
  LMState* st = h->d_state.as<LMState>();
  EnqueueEval(h);
  const int rc = EnqueueLinearise(h, o, false);
  LAUNCH(h, ctrl_after_jacobian_kernel, 1, 1, st, h->d_red.as<double>(), recs, o);
  return rc;
}
int EnqueuePass(bsnls_handle* h, const Options& o, bsnls_iteration* recs) {   // one LM iteration
  LMState* st = h->d_state.as<LMState>();
  EnqueueEval(h, 1);
  if (!h->fused) EnqueueLandmark(h, o);
  int rc = EnqueueLinearise(h, o, true);
  if (rc != BSNLS_OK) return rc;
  LAUNCH(h, ctrl_after_jacobian_kernel, 1, 1, st, h->d_red.as<double>(), recs, o);
  EnqueueAssembleFinish(h, o);
  if (h->pcg) { rc = EnqueueSparseSolve(h); if (rc != BSNLS_OK) return rc; } else EnqueueDenseSolve(h);
  if ((rc = EnqueueStep(h)) != BSNLS_OK) return rc;
  LAUNCH(h, ctrl_end_iteration_kernel, 1, 1, st, h->d_red.as<double>(), recs, o);
  LAUNCH(h, apply_step_kernel, BlocksFor((int)h->num_values, 256), 256, h->d_x.as<double>(), h->d_xc.as<double>(), (int)h->num_values, st);
  return BSNLS_OK;
}

}  // namespace

extern "C" {

void bsnls_default_options(bsnls_options* o) {
  o->max_num_iterations = 50; o->max_solver_time_in_seconds = 1e9; o->function_tolerance = 1e-6;
  o->gradient_tolerance = 1e-10; o->parameter_tolerance = 1e-8; o->initial_trust_region_radius = 1e4;
  o->max_trust_region_radius = 1e16; o->min_trust_region_radius = 1e-32; o->min_relative_decrease = 1e-3;
  o->min_lm_diagonal = 1e-6; o->max_lm_diagonal = 1e32; o->max_num_consecutive_invalid_steps = 5;
  o->jacobi_scaling = 1; o->linear_solver = BSNLS_LINEAR_AUTO; o->pcg_max_iterations = 500; o->pcg_tolerance = 1e-13;
  o->device = -1;
}

const char* bsnls_last_error(const bsnls_handle* h) { return h ? h->error.c_str() : g_create_error.c_str(); }
int64_t bsnls_launch_count(const bsnls_handle* h) { return h ? h->launches : 0; }

int bsnls_create(const bsnls_options* options, bsnls_handle** out) {
  if (!out) return BSNLS_ERR_INVALID;
  *out = nullptr;
  bsnls_options o; if (options) o = *options; else bsnls_default_options(&o);
  std::string why;
  if (!ValidOptions(o, &why)) { g_create_error = "invalid options: " + why; return BSNLS_ERR_INVALID; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU path)";
    return BSNLS_ERR_CUDA;
  }
  int dev = o.device;
  if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) dev = 0; }
  if (dev >= ndev) { g_create_error = "device ordinal out of range"; return BSNLS_ERR_INVALID; }
  if ((e = cudaSetDevice(dev)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return BSNLS_ERR_CUDA; }
  auto* h = new bsnls_handle();
  h->opt = o; h->device = dev;
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreate(&h->ev0)) != cudaSuccess || (e = cudaEventCreate(&h->ev1)) != cudaSuccess ||
      (e = cudaEventCreate(&h->ev2)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&h->ev_stage, cudaEventDisableTiming)) != cudaSuccess || (e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)) != cudaSuccess || (e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming)) != cudaSuccess) {
    g_create_error = cudaGetErrorString(e); delete h; return BSNLS_ERR_CUDA;
  }
  if ((e = cudaFuncSetAttribute(chunk_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, syrk_smem_bytes(kChunkSlots))) != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete h; return BSNLS_ERR_CUDA; }
  if ((e = cudaFuncSetAttribute(chunk_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fused_smem_bytes(kChunkSlots))) != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete h; return BSNLS_ERR_CUDA; }
  if ((e = cudaFuncSetAttribute(tile_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(double) * kPanelSmemDoubles))) != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete h; return BSNLS_ERR_CUDA; }
  if ((e = cudaFuncSetAttribute(tile_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(double) * kPanelSmemDoubles))) != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete h; return BSNLS_ERR_CUDA; }
  for (auto& ev : h->pev) if ((e = cudaEventCreate(&ev)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete h; return BSNLS_ERR_CUDA; }
  *out = h;
  return BSNLS_OK;
}

void bsnls_destroy(bsnls_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  if (h->graph_pass) cudaGraphExecDestroy(h->graph_pass);
  if (h->graph_tail) cudaGraphExecDestroy(h->graph_tail);
  if (h->graph_rp) cudaGraphExecDestroy(h->graph_rp);
  if (h->comm) Nccl().CommDestroy(h->comm);
  h->arena.Release();
  if (h->pcgd_done_host) cudaFreeHost(h->pcgd_done_host);
  if (h->stream) cudaStreamDestroy(h->stream);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->ev_stage) cudaEventDestroy(h->ev_stage);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev2) cudaEventDestroy(h->ev2);
  for (auto& ev : h->pev) if (ev) cudaEventDestroy(ev);
  delete h;
}

int bsnls_set_options(bsnls_handle* h, const bsnls_options* o) {
  if (!h || !o) return BSNLS_ERR_INVALID;
  std::string why;
  if (!ValidOptions(*o, &why)) return h->fail(BSNLS_ERR_INVALID, "invalid options: %s", why.c_str());
  const int dev = h->opt.device;
  bsnls_options n = *o; n.device = dev;
  const bool solver_changed = n.linear_solver != h->opt.linear_solver;
  // kernel arguments (options) are baked into the captured graphs: re-capture only when something changed
  { const bsnls_options &a = n, &b = h->opt;   // field by field (the struct has padding); the time budget is a host-side check
    const bool same = a.max_num_iterations == b.max_num_iterations && a.function_tolerance == b.function_tolerance &&
        a.gradient_tolerance == b.gradient_tolerance && a.parameter_tolerance == b.parameter_tolerance &&
        a.initial_trust_region_radius == b.initial_trust_region_radius && a.max_trust_region_radius == b.max_trust_region_radius &&
        a.min_trust_region_radius == b.min_trust_region_radius && a.min_relative_decrease == b.min_relative_decrease &&
        a.min_lm_diagonal == b.min_lm_diagonal && a.max_lm_diagonal == b.max_lm_diagonal &&
        a.max_num_consecutive_invalid_steps == b.max_num_consecutive_invalid_steps && a.jacobi_scaling == b.jacobi_scaling &&
        a.linear_solver == b.linear_solver && a.pcg_max_iterations == b.pcg_max_iterations && a.pcg_tolerance == b.pcg_tolerance;
    if (!same) { h->graphs_valid = false; h->graph_rp_valid = false; } }
  h->opt = n;
  if (solver_changed && h->has_problem) h->has_problem = false;   // the reduced-system storage depends on the linear solver: set_problem again
  return BSNLS_OK;
}

int bsnls_set_problem(bsnls_handle* h, const bsnls_problem* pr) {
  NvtxRange nvtx_range("bsnls_set_problem");
  if (!h || !pr) return BSNLS_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  h->has_problem = false; h->jac_valid = false;
  const bool dbg = std::getenv("BSNLS_DEBUG") != nullptr;
  auto tphase = std::chrono::steady_clock::now();
  auto phase = [&](const char* name) {   // BSNLS_DEBUG: host time of each analysis phase
    if (!dbg) return;
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[bsnls] set_problem %-28s %8.3f ms\n", name, 1e3 * std::chrono::duration<double>(now - tphase).count());
    tphase = now;
  };
  const int nb = pr->num_blocks;
  if (nb <= 0) return h->fail(BSNLS_ERR_INVALID, "problem has no parameter blocks");
  // ---- blocks ----
  h->num_blocks = nb; h->blk_off.resize(nb); h->blk_size.assign(pr->block_size, pr->block_size + nb);
  h->blk_manifold.assign(pr->block_manifold, pr->block_manifold + nb);
  int64_t off = 0;
  for (int b = 0; b < nb; ++b) {
    h->blk_off[b] = (int)off; off += pr->block_size[b];
    const int sz = pr->block_size[b], mf = pr->block_manifold[b];
    if (!((mf == BSNLS_MANIFOLD_EUCLIDEAN && sz == 3) || (mf == BSNLS_MANIFOLD_QUATERNION && sz == 4)))
      if (!pr->block_constant[b])
        return h->fail(BSNLS_ERR_UNSUPPORTED, "block %d: only 3-vectors and quaternions are supported (size %d, manifold %d)", b, sz, mf);
  }
  if (off > 0x7fffffffLL) return h->fail(BSNLS_ERR_UNSUPPORTED, "too many parameters");
  h->num_values = off;
  // ---- residual blocks in program order ----
  int64_t total = 0;
  for (int g = 0; g < pr->num_groups; ++g) {
    if (pr->groups[g].type < 0 || pr->groups[g].type >= BSNLS_NUM_FACTOR_TYPES) return h->fail(BSNLS_ERR_UNSUPPORTED, "unknown factor type %d", pr->groups[g].type);
    total += pr->groups[g].count;
  }
  // active blocks (ceres reduced program: constant blocks and unreferenced blocks drop out)
  std::vector<char> active(nb, 0);
  const bsnls_factor_group* rg = nullptr;
  std::vector<const bsnls_factor_group*> gfg;
  for (int g = 0; g < pr->num_groups; ++g) {
    const bsnls_factor_group& G = pr->groups[g];
    if (G.type == BSNLS_FACTOR_REPROJECTION) { if (rg) return h->fail(BSNLS_ERR_INVALID, "two reprojection groups"); rg = &G; }
    else gfg.push_back(&G);
    const int nbk = kNumBlocks[G.type];
    for (int64_t i = 0; i < G.count; ++i) {
      bool any = false;
      const int used = UsedBlocks(G, i);
      if (G.type == BSNLS_FACTOR_MARGINAL) {   // the shape in consts must agree with the block list and the block sizes
        const double* c = G.consts + i * BSNLS_MARGINAL_STRIDE;
        if (used < 1 || (int)c[0] != used) return h->fail(BSNLS_ERR_INVALID, "marginal factor: block count and consts[0] disagree");
        for (int b = 0; b < used; ++b) { const int blk = G.blocks[i * nbk + b]; if (blk >= 0 && blk < nb && (int)c[1 + b] != pr->block_size[blk]) return h->fail(BSNLS_ERR_INVALID, "marginal factor: block size table disagrees with the problem"); }
      }
      for (int b = 0; b < used; ++b) {
        const int blk = G.blocks[i * nbk + b];
        if (blk < 0 || blk >= nb) return h->fail(BSNLS_ERR_INVALID, "block index out of range");
        any = any || !pr->block_constant[blk];
      }
      if (any) for (int b = 0; b < used; ++b) { const int blk = G.blocks[i * nbk + b]; if (!pr->block_constant[blk]) active[blk] = 1; }
    }
  }
  std::sort(gfg.begin(), gfg.end(), [](const bsnls_factor_group* a, const bsnls_factor_group* b) { return a->type < b->type; });
  h->eblocks.clear(); h->fblocks.clear(); h->active_blocks.clear();
  std::vector<int> epos(nb, -1), fpos(nb, -1);
  for (int b = 0; b < nb; ++b) {
    if (!active[b]) continue;
    h->active_blocks.push_back(b);
    if (pr->block_group[b] == 0) { epos[b] = (int)h->eblocks.size(); h->eblocks.push_back(b); }
    else { fpos[b] = (int)h->fblocks.size(); h->fblocks.push_back(b); }
  }
  const int ne_g = (int)h->eblocks.size();   // eliminated blocks of the whole graph; this rank keeps the set `mine` of them
  h->ne_global = ne_g; h->nf = (int)h->fblocks.size();
  const int nf_dim = 3 * h->nf;
  if (nf_dim == 0) return h->fail(BSNLS_ERR_UNSUPPORTED, "no non-landmark parameter blocks to solve for");
  // assembly layout and linear solver: small windows dense + dense; mid-size block-sparse assembly expanded into the
  // dense Cholesky; large (or on request) block-sparse + PCG
  // assembly layout and linear solver: sliding windows (<= kRbMaxN unknowns) dense storage + a single-CTA Cholesky; everything
  // larger is assembled block-sparse and factored by the tile-sparse Cholesky in a bandwidth-reducing order, unless the
  // iterative solver is requested or the fill of the tile factor is too large (decided after the symbolic step below)
  h->pcg = h->opt.linear_solver == BSNLS_LINEAR_ITERATIVE_SCHUR;
  h->tiled = !h->pcg && nf_dim > kRbMaxN;
  h->sparse = h->pcg || h->tiled;

  // ---- reprojection: poses, landmark-sorted observations ----
  std::vector<int> pose_of_q(nb, -1), pose_of_t(nb, -1);
  h->pose_q_blk.clear(); h->pose_t_blk.clear();
  const int nobs = rg ? rg->count : 0;
  std::vector<int> obs_pose(nobs), obs_lm(nobs), order(nobs);
  std::vector<int> lm_count(ne_g + 1, 0);
  for (int i = 0; i < nobs; ++i) {
    const int q = rg->blocks[3 * (int64_t)i], t = rg->blocks[3 * (int64_t)i + 1], p = rg->blocks[3 * (int64_t)i + 2];
    if (epos[p] < 0) return h->fail(BSNLS_ERR_UNSUPPORTED, "reprojection factor %d: landmark block must be non-constant and in elimination group 0", i);
    if (epos[q] >= 0 || epos[t] >= 0) return h->fail(BSNLS_ERR_UNSUPPORTED, "reprojection factor %d: pose blocks must not be in elimination group 0", i);
    if (pr->block_size[q] != 4 || pr->block_size[t] != 3 || pr->block_size[p] != 3) return h->fail(BSNLS_ERR_INVALID, "reprojection factor %d: block sizes must be 4,3,3", i);
    int pose = pose_of_q[q];
    if (pose < 0 && pose_of_t[t] < 0) { pose = (int)h->pose_q_blk.size(); h->pose_q_blk.push_back(q); h->pose_t_blk.push_back(t); pose_of_q[q] = pose; pose_of_t[t] = pose; }
    else if (pose < 0 || pose_of_t[t] != pose) return h->fail(BSNLS_ERR_UNSUPPORTED, "orientation/position blocks are paired inconsistently across reprojection factors");
    if (rg->aux[i] < 0 || rg->aux[i] >= pr->num_cameras) return h->fail(BSNLS_ERR_INVALID, "camera index out of range");
    obs_pose[i] = pose; obs_lm[i] = epos[p]; lm_count[epos[p] + 1]++;
  }
  h->np = (int)h->pose_q_blk.size(); h->nobs_global = nobs;
  for (int e = 0; e < ne_g; ++e) { if (lm_count[e + 1] == 0) return h->fail(BSNLS_ERR_UNSUPPORTED, "group-0 block without reprojection factors"); lm_count[e + 1] += lm_count[e]; }
  std::vector<int> lm_ptr = lm_count;
  { std::vector<int> cur(lm_ptr.begin(), lm_ptr.end() - 1); for (int i = 0; i < nobs; ++i) order[cur[obs_lm[i]]++] = i; }  // stable
  std::vector<int> g_pose(nobs);   // pose of every observation of the graph, landmark-sorted (the pair pattern is global)
  for (int k = 0; k < nobs; ++k) g_pose[k] = obs_pose[order[k]];

  phase("blocks + observation sort");
  // ---- pose-pair blocks of the reduced system: pattern of the WHOLE graph (identical on every rank) ----
  // Pair (a <= b) exists iff some landmark is seen by both poses. Built per pose with a stamp array over the poses of
  // the landmarks it observes (no hashing); pair id = position in (a, b) order; lookups go through pair_row / pair_b.
  std::vector<int> pair_row;            // CSR over a: pairs [pair_row[a], pair_row[a+1]) have pose a and ascending b >= a
  h->pair_a.clear(); h->pair_b.clear();
  if (h->np <= 64) {
    // sliding windows: one 64-bit pose mask per landmark, OR-ed into the rows of the poses it contains
    std::vector<uint64_t> row(h->np, 0);
    for (int e = 0; e < ne_g; ++e) {
      uint64_t m = 0;
      for (int k = lm_ptr[e]; k < lm_ptr[e + 1]; ++k) m |= uint64_t(1) << g_pose[k];
      for (uint64_t t = m; t; t &= t - 1) row[__builtin_ctzll(t)] |= m;
    }
    pair_row.assign(h->np + 1, 0);
    for (int a = 0; a < h->np; ++a) {
      for (uint64_t t = row[a] >> a; t; t &= t - 1) { h->pair_a.push_back(a); h->pair_b.push_back(a + __builtin_ctzll(t)); }
      pair_row[a + 1] = (int)h->pair_b.size();
    }
    h->npairs = (int)h->pair_b.size();
  } else
  {
    std::vector<int> gp_ptr(h->np + 1, 0), gp_lm(nobs);
    for (int k = 0; k < nobs; ++k) gp_ptr[g_pose[k] + 1]++;
    for (int a = 0; a < h->np; ++a) gp_ptr[a + 1] += gp_ptr[a];
    { std::vector<int> cur(gp_ptr.begin(), gp_ptr.end() - 1); for (int e = 0; e < ne_g; ++e) for (int k = lm_ptr[e]; k < lm_ptr[e + 1]; ++k) gp_lm[cur[g_pose[k]]++] = e; }
    // rows of different poses are independent: host threads take slices of the poses, each with its own stamp array
    std::vector<std::vector<int>> rows(h->np);
    ParallelFor(h->np, 64, [&](int a0, int a1) {
      std::vector<int> stamp(h->np, -1);
      for (int a = a0; a < a1; ++a) {
        std::vector<int>& row = rows[a];
        for (int q = gp_ptr[a]; q < gp_ptr[a + 1]; ++q) {
          const int e = gp_lm[q];
          for (int k = lm_ptr[e]; k < lm_ptr[e + 1]; ++k) { const int b = g_pose[k]; if (b >= a && stamp[b] != a) { stamp[b] = a; row.push_back(b); } }
        }
        std::sort(row.begin(), row.end());
      }
    });
    pair_row.assign(h->np + 1, 0);
    for (int a = 0; a < h->np; ++a) {
      for (int b : rows[a]) { h->pair_a.push_back(a); h->pair_b.push_back(b); }
      pair_row[a + 1] = (int)h->pair_b.size();
    }
    h->npairs = (int)h->pair_b.size();
  }
  std::vector<int> pid_dense;   // small pose counts: direct table instead of a search per lookup
  if (h->np <= 256) {
    pid_dense.assign((size_t)h->np * h->np, -1);
    for (int k = 0; k < h->npairs; ++k) { pid_dense[(size_t)h->pair_a[k] * h->np + h->pair_b[k]] = k; pid_dense[(size_t)h->pair_b[k] * h->np + h->pair_a[k]] = k; }
  }
  auto pair_id = [&](int a2, int b2) -> int {   // -1: the two poses share no landmark
    if (!pid_dense.empty()) return pid_dense[(size_t)a2 * h->np + b2];
    const int a = std::min(a2, b2), b = std::max(a2, b2);
    const int* r0 = h->pair_b.data() + pair_row[a]; const int* r1 = h->pair_b.data() + pair_row[a + 1];
    const int* p = std::lower_bound(r0, r1, b);
    return (p != r1 && *p == b) ? (int)(p - h->pair_b.data()) : -1;
  };
  phase("pair pattern");
  // ---- pose ranking for locality: breadth-first order of the co-visibility graph from a peripheral pose ----
  std::vector<int> pose_rank(h->np, -1);
  {
    std::vector<std::vector<int>> adj(h->np);
    for (int p2 = 0; p2 < h->npairs; ++p2) if (h->pair_a[p2] != h->pair_b[p2]) { adj[h->pair_a[p2]].push_back(h->pair_b[p2]); adj[h->pair_b[p2]].push_back(h->pair_a[p2]); }
    auto bfs = [&](int start, std::vector<int>& out) {
      std::vector<char> seen(h->np, 0); out.clear(); out.push_back(start); seen[start] = 1;
      for (size_t q = 0; q < out.size(); ++q) {
        std::vector<int> nb2; for (int v : adj[out[q]]) if (!seen[v]) { seen[v] = 1; nb2.push_back(v); }
        std::sort(nb2.begin(), nb2.end(), [&](int x, int y2) { return adj[x].size() != adj[y2].size() ? adj[x].size() < adj[y2].size() : x < y2; });
        out.insert(out.end(), nb2.begin(), nb2.end());
      }
    };
    int next_rank = 0; std::vector<int> comp, comp2;
    for (int p0 = 0; p0 < h->np; ++p0) {
      if (pose_rank[p0] >= 0) continue;
      bfs(p0, comp); bfs(comp.back(), comp2);      // second sweep starts at a far end of the component
      for (int v : comp2) pose_rank[v] = next_rank++;
    }
  }
  // ---- locality order of the landmarks: by the earliest (then the latest) pose that sees them, poses ranked along the
  // trajectory (breadth-first order above). Consecutive landmarks then share poses, which keeps chunks tight; and the
  // shard of a rank is a contiguous range of THIS order balanced by observation count, so a rank touches the poses of
  // one stretch of the trajectory (full Jacobian-pass CTAs, few pair blocks) instead of every pose of the graph. ----
  std::vector<int> mine;                           // this rank's landmarks: positions in the elimination order, in locality order
  std::vector<char> is_mine(ne_g, 0);
  {
    std::vector<std::pair<int, int>> key(ne_g);
    std::vector<int> gl(ne_g);
    for (int e = 0; e < ne_g; ++e) {
      int mn = 1 << 30, mx = -1;
      for (int k = lm_ptr[e]; k < lm_ptr[e + 1]; ++k) { mn = std::min(mn, pose_rank[g_pose[k]]); mx = std::max(mx, pose_rank[g_pose[k]]); }
      key[e] = {mn, mx}; gl[e] = e;
    }
    std::stable_sort(gl.begin(), gl.end(), [&](int x, int y2) { return key[x] < key[y2]; });
    std::vector<int> prefix(ne_g + 1, 0);
    for (int q = 0; q < ne_g; ++q) prefix[q + 1] = prefix[q] + (lm_ptr[gl[q] + 1] - lm_ptr[gl[q]]);
    int lo = 0, hi = ne_g;
    bsnls_shard_range(prefix.data(), ne_g, h->rank, h->world, &lo, &hi);
    mine.assign(gl.begin() + lo, gl.begin() + hi);
    for (int e : mine) is_mine[e] = 1;
    h->nobs = prefix[hi] - prefix[lo];
  }
  h->ne = (int)mine.size();
  h->shard_e.assign(mine.begin(), mine.end());
  const int ne_dim = 3 * h->ne;
  const int nobs_local = h->nobs;
  // ---- internal order of the f-blocks. The API order (h->fblocks, program order) keeps the two blocks of a pose far
  // apart, which makes the reduced system's envelope nearly full. For the tile-sparse factorisation the blocks are
  // renumbered by a nested dissection of the block graph (pose pairs + generic-factor cliques). From here on fpos[] is
  // the INTERNAL position; f_api[] maps back.
  h->f_api.resize(h->nf); for (int k = 0; k < h->nf; ++k) h->f_api[k] = k;
  if (h->tiled) {
    std::vector<std::vector<int>> adj(h->nf);
    auto link = [&](int x, int y2) { if (x >= 0 && y2 >= 0 && x != y2) { adj[x].push_back(y2); adj[y2].push_back(x); } };
    for (int p2 = 0; p2 < h->npairs; ++p2) {
      const int a = h->pair_a[p2], b = h->pair_b[p2];
      const int va[2] = {fpos[h->pose_q_blk[a]], fpos[h->pose_t_blk[a]]}, vb[2] = {fpos[h->pose_q_blk[b]], fpos[h->pose_t_blk[b]]};
      if (a == b) link(va[0], va[1]); else for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) link(va[i], vb[j]);
    }
    for (const bsnls_factor_group* G : gfg) {
      const int nbk = kNumBlocks[G->type];
      for (int64_t i = 0; i < G->count; ++i) {
        const int used = UsedBlocks(*G, i);
        for (int b1 = 0; b1 < used; ++b1) for (int b2 = 0; b2 < b1; ++b2) link(fpos[G->blocks[i * nbk + b1]], fpos[G->blocks[i * nbk + b2]]);
      }
    }
    for (auto& v : adj) { std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end()); }
    // Two candidate orders, each with the slot (position in the padded sequence, kTileBlocks per tile) of every block; the
    // one whose tile-level symbolic factorisation is cheaper wins (BSNLS_ORDER=graph|band forces one).
    struct Candidate { std::vector<int> order, slot; int nT = 0, levels = 0; long long tiles = 0, products = 0; double cost = 0.0; };
    int nd_leaf = kNdLeaf;                          // BSNLS_ND_LEAF: tuning knob (tiles per leaf of the dissection)
    if (const char* ev = std::getenv("BSNLS_ND_LEAF")) nd_leaf = std::max(1, std::atoi(ev));
    auto evaluate = [&](Candidate& c) {             // levels of the elimination tree and tile products, as the factorisation below finds them
      std::vector<int> tile(h->nf); for (int k = 0; k < h->nf; ++k) tile[c.order[k]] = c.slot[k] / kTileBlocks;
      std::vector<std::vector<int>> col(c.nT);
      for (int v = 0; v < h->nf; ++v) for (int u : adj[v]) if (tile[u] > tile[v]) col[tile[v]].push_back(tile[u]);
      std::vector<int> level(c.nT, 0); c.levels = 0; c.tiles = 0; c.products = 0;
      for (int k = 0; k < c.nT; ++k) {
        std::sort(col[k].begin(), col[k].end()); col[k].erase(std::unique(col[k].begin(), col[k].end()), col[k].end());
        const long long a = (long long)col[k].size(); c.tiles += 1 + a; c.products += a * (a + 1) / 2; c.levels = std::max(c.levels, level[k] + 1);
        if (a > 0) { const int par = col[k][0]; col[par].insert(col[par].end(), col[k].begin() + 1, col[k].end()); level[par] = std::max(level[par], level[k] + 1); }
      }
      c.cost = 30.0 * c.levels + 0.03 * (double)c.products;   // microseconds: a launch pair per level is latency, the products are throughput
    };
    // (a) Nested dissection of the block graph (George's automatic dissection: the separator is a level set of a
    // breadth-first sweep from a pseudo-peripheral block, picked small and near the middle; blocks of it without a
    // neighbour on the far side go back to the near side). Done on blocks, not tiles: a loop closure costs the separator
    // one 3-block, not a whole tile, so a trajectory with closures every few keyframes keeps separators of a tile or two.
    // Every region (leaf or separator) then starts on a tile boundary -- the slots left over are identity padding -- so
    // that tile columns of different leaves stay independent.
    auto order_graph = [&](Candidate& c) {
      std::vector<int>& orderv = c.order; orderv.clear(); orderv.reserve(h->nf);
      std::vector<int> region_end;
      const int leaf_blocks = nd_leaf * kTileBlocks;
      std::vector<int> in_set(h->nf, -1), lev(h->nf, 0), vis(h->nf, 0); int vis_id = 0, next_id = 0;
      auto bfs = [&](int start, int id, std::vector<int>& out) {   // Cuthill-McKee sweep over the blocks of set `id`
        ++vis_id; out.clear(); out.push_back(start); vis[start] = vis_id; lev[start] = 0;
        for (size_t q = 0; q < out.size(); ++q) {
          const int u = out[q]; const size_t first = out.size();
          for (int v : adj[u]) if (in_set[v] == id && vis[v] != vis_id) { vis[v] = vis_id; lev[v] = lev[u] + 1; out.push_back(v); }
          std::sort(out.begin() + first, out.end(), [&](int x, int y2) { return adj[x].size() != adj[y2].size() ? adj[x].size() < adj[y2].size() : x < y2; });
        }
      };
      std::function<void(std::vector<int>&)> nd = [&](std::vector<int>& nodes) {
        if (nodes.empty()) return;
        const int id = next_id++;
        for (int v : nodes) in_set[v] = id;
        std::vector<int> o1, o2;
        bfs(nodes[0], id, o1);
        if (o1.size() < nodes.size()) {               // not connected: one component at a time
          std::vector<int> rest; for (int v : nodes) if (vis[v] != vis_id) rest.push_back(v);
          nd(o1); nd(rest); return;
        }
        bfs(o1.back(), id, o2); bfs(o2.back(), id, o1);
        const int L = lev[o1.back()];
        if ((int)nodes.size() <= leaf_blocks || L < 2) {
          orderv.insert(orderv.end(), o1.rbegin(), o1.rend()); region_end.push_back((int)orderv.size()); return;
        }
        std::vector<int> cnt(L + 1, 0); for (int v : o1) cnt[lev[v]]++;
        int best = 1; long long best_score = -1, below = cnt[0];
        for (int l = 1; l < L; ++l) {
          const long long above = (long long)o1.size() - below - cnt[l], score = 4LL * cnt[l] + std::llabs(below - above);
          if (best_score < 0 || score < best_score) { best_score = score; best = l; }
          below += cnt[l];
        }
        std::vector<int> A, B, S;
        for (int v : o1) {
          if (lev[v] < best) A.push_back(v);
          else if (lev[v] > best) B.push_back(v);
          else { bool far = false; for (int u : adj[v]) if (in_set[u] == id && lev[u] == best + 1) { far = true; break; }
                 (far ? S : A).push_back(v); }
        }
        // leaves give the blocks beyond their last full tile to the separator while its last tile has room (A ends and B
        // starts next to the separator): a tile less per leaf at no cost to the separator
        int room = (kTileBlocks - (int)S.size() % kTileBlocks) % kTileBlocks;
        if ((int)A.size() <= leaf_blocks && (int)A.size() > kTileBlocks) { const int r = (int)A.size() % kTileBlocks; if (r > 0 && r <= room) { S.insert(S.end(), A.end() - r, A.end()); A.resize(A.size() - r); room -= r; } }
        if ((int)B.size() <= leaf_blocks && (int)B.size() > kTileBlocks) { const int r = (int)B.size() % kTileBlocks; if (r > 0 && r <= room) { S.insert(S.end(), B.begin(), B.begin() + r); B.erase(B.begin(), B.begin() + r); room -= r; } }
        nd(A); nd(B);
        orderv.insert(orderv.end(), S.begin(), S.end()); region_end.push_back((int)orderv.size());
      };
      std::vector<int> all(h->nf); for (int k = 0; k < h->nf; ++k) all[k] = k;
      nd(all);
      c.slot.assign(h->nf, 0); int slot = 0, k0 = 0;
      for (int e : region_end) { for (int k = k0; k < e; ++k) c.slot[k] = slot + (k - k0); slot += (e - k0 + kTileBlocks - 1) / kTileBlocks * kTileBlocks; k0 = e; }
      c.nT = slot / kTileBlocks;
    };
    // (b) Reverse Cuthill-McKee line (a trajectory becomes a band) cut by a nested dissection at tile granularity: the
    // line of tiles is cut in the middle by a separator as wide as the band (plus one endpoint tile of every long edge
    // that crosses the cut), the halves are ordered recursively and the separator goes last. No padding: the better
    // choice for a pure band, the worse one as soon as long edges are frequent.
    auto order_band = [&](Candidate& c) {
      std::vector<int>& orderv = c.order; orderv.clear(); orderv.reserve(h->nf);
      std::vector<char> seen(h->nf, 0);
      auto bfs = [&](int start, std::vector<int>& out, bool mark) {
        std::vector<char> local; std::vector<char>& sn = mark ? seen : local; if (!mark) local.assign(h->nf, 0);
        const size_t first = out.size(); out.push_back(start); sn[start] = 1;
        for (size_t q = first; q < out.size(); ++q) {
          std::vector<int> nb2; for (int v : adj[out[q]]) if (!sn[v] && (mark || !seen[v])) { sn[v] = 1; nb2.push_back(v); }
          std::sort(nb2.begin(), nb2.end(), [&](int x, int y2) { return adj[x].size() != adj[y2].size() ? adj[x].size() < adj[y2].size() : x < y2; });
          out.insert(out.end(), nb2.begin(), nb2.end());
        }
      };
      for (int v0 = 0; v0 < h->nf; ++v0) {
        if (seen[v0]) continue;
        std::vector<int> probe; bfs(v0, probe, false);           // pseudo-peripheral start: the far end of a sweep from v0
        std::vector<int> probe2; bfs(probe.back(), probe2, false);
        bfs(probe2.back(), orderv, true);
      }
      std::reverse(orderv.begin(), orderv.end());
      const int nTf = h->nf / kTileBlocks;            // full tiles; a partial last tile stays last
      if (nTf > nd_leaf) {
        std::vector<int> pos(h->nf); for (int k = 0; k < h->nf; ++k) pos[orderv[k]] = k;
        std::vector<std::vector<int>> tadj(nTf);
        for (int u = 0; u < h->nf; ++u) {
          const int tu = pos[u] / kTileBlocks; if (tu >= nTf) continue;
          for (int v : adj[u]) { const int tv = pos[v] / kTileBlocks; if (tv < nTf && tv != tu) tadj[tu].push_back(tv); }
        }
        for (auto& v : tadj) { std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end()); }
        std::vector<int> torder, idx(nTf, -1); torder.reserve(nTf);
        std::function<void(std::vector<int>&)> nd = [&](std::vector<int>& tl) {
          const int m = (int)tl.size();
          if (m <= nd_leaf) { torder.insert(torder.end(), tl.begin(), tl.end()); return; }
          for (int a = 0; a < m; ++a) idx[tl[a]] = a;
          const int mid = m / 2; int w = 0;
          for (int a = 0; a < mid; ++a) for (int v : tadj[tl[a]]) { const int b = idx[v]; if (b >= mid && b - a <= kNdShort) w = std::max(w, b - mid + 1); }
          std::vector<char> insep(m, 0);
          for (int a = mid; a < std::min(m, mid + w); ++a) insep[a] = 1;
          for (int a = 0; a < mid; ++a) for (int v : tadj[tl[a]]) { const int b = idx[v]; if (b >= mid + w) insep[b] = 1; }
          for (int a = 0; a < m; ++a) idx[tl[a]] = -1;
          std::vector<int> left(tl.begin(), tl.begin() + mid), right, sep;
          for (int a = mid; a < m; ++a) (insep[a] ? sep : right).push_back(tl[a]);
          nd(left); nd(right); torder.insert(torder.end(), sep.begin(), sep.end());
        };
        std::vector<int> all(nTf); for (int t = 0; t < nTf; ++t) all[t] = t;
        nd(all);
        std::vector<int> order2; order2.reserve(h->nf);
        for (int tp = 0; tp < nTf; ++tp) for (int j = 0; j < kTileBlocks; ++j) order2.push_back(orderv[torder[tp] * kTileBlocks + j]);
        for (int k = nTf * kTileBlocks; k < h->nf; ++k) order2.push_back(orderv[k]);
        orderv.swap(order2);
      }
      c.slot.resize(h->nf); for (int k = 0; k < h->nf; ++k) c.slot[k] = k;
      c.nT = (h->nf + kTileBlocks - 1) / kTileBlocks;
    };
    Candidate cg, cb;
    const char* force = std::getenv("BSNLS_ORDER");
    const bool want_graph = !force || std::strcmp(force, "band") != 0, want_band = !force || std::strcmp(force, "graph") != 0;
    if (want_graph) { order_graph(cg); evaluate(cg); }
    if (want_band) { order_band(cb); evaluate(cb); }
    const bool pick_graph = want_graph && (!want_band || cg.cost < cb.cost);
    if (std::getenv("BSNLS_DEBUG")) {
      if (want_graph) std::fprintf(stderr, "[bsnls] order 'graph': %d tile columns, %d levels, %lld tiles, %lld products (cost %.0f)%s\n", cg.nT, cg.levels, cg.tiles, cg.products, cg.cost, pick_graph ? " <- used" : "");
      if (want_band) std::fprintf(stderr, "[bsnls] order 'band':  %d tile columns, %d levels, %lld tiles, %lld products (cost %.0f)%s\n", cb.nT, cb.levels, cb.tiles, cb.products, cb.cost, pick_graph ? "" : " <- used");
    }
    Candidate& win = pick_graph ? cg : cb;
    std::vector<int> rank(h->nf, -1); std::vector<int>& orderv = win.order;
    h->tc_blk_slot = win.slot; h->tc_nT = win.nT;
    for (int k = 0; k < h->nf; ++k) { rank[orderv[k]] = k; h->f_api[k] = orderv[k]; }
    for (int b = 0; b < nb; ++b) if (fpos[b] >= 0) fpos[b] = rank[fpos[b]];
  }
  phase("pose / f-block ordering");
  h->blk_tan.assign(nb, -1);
  for (int b = 0; b < nb; ++b) if (fpos[b] >= 0) h->blk_tan[b] = ne_dim + 3 * fpos[b];
  for (int e = 0; e < h->ne; ++e) h->blk_tan[h->eblocks[mine[e]]] = 3 * e;
  // active list: local landmarks, padded to a CTA boundary, then the (replicated) f-blocks
  h->active_blocks.clear();
  for (int e = 0; e < h->ne; ++e) h->active_blocks.push_back(h->eblocks[mine[e]]);
  while (h->active_blocks.size() % 256) h->active_blocks.push_back(-1);
  h->nb_act_e = (int)h->active_blocks.size() / 256;
  for (int b : h->fblocks) h->active_blocks.push_back(b);
  std::vector<int> s_pose(nobs_local), s_lm(nobs_local), s_cam(nobs_local), inv_order(nobs, -1), lm_ptr_local(h->ne + 1, 0), lm_off(h->ne);
  std::vector<double> s_pixel(2 * (size_t)nobs_local);
  {
    for (int e = 0; e < h->ne; ++e) lm_ptr_local[e + 1] = lm_ptr_local[e] + (lm_ptr[mine[e] + 1] - lm_ptr[mine[e]]);
    ParallelFor(h->ne, 20000, [&](int e0, int e1) {
      for (int e = e0; e < e1; ++e) {
        const int ge = mine[e];
        lm_off[e] = h->blk_off[h->eblocks[ge]];
        int kl = lm_ptr_local[e];
        for (int k = lm_ptr[ge]; k < lm_ptr[ge + 1]; ++k, ++kl) {
          const int i = order[k]; inv_order[i] = kl;
          s_pose[kl] = obs_pose[i]; s_lm[kl] = e; s_cam[kl] = rg->aux[i];
          s_pixel[2 * (size_t)kl] = rg->consts[2 * (size_t)i]; s_pixel[2 * (size_t)kl + 1] = rg->consts[2 * (size_t)i + 1];
        }
      }
    });
  }
  std::vector<int> pose_q_off(h->np), pose_t_off(h->np), pose_q_tan(h->np), pose_t_tan(h->np), pose_ptr(h->np + 1, 0), pose_obs(nobs_local);
  for (int a = 0; a < h->np; ++a) {
    pose_q_off[a] = h->blk_off[h->pose_q_blk[a]]; pose_t_off[a] = h->blk_off[h->pose_t_blk[a]];
    pose_q_tan[a] = h->blk_tan[h->pose_q_blk[a]]; pose_t_tan[a] = h->blk_tan[h->pose_t_blk[a]];
  }
  for (int k = 0; k < nobs_local; ++k) pose_ptr[s_pose[k] + 1]++;
  for (int a = 0; a < h->np; ++a) pose_ptr[a + 1] += pose_ptr[a];
  { std::vector<int> cur(pose_ptr.begin(), pose_ptr.end() - 1); for (int k = 0; k < nobs_local; ++k) pose_obs[cur[s_pose[k]]++] = k; }
  // pose-major evaluation items: <= kEvalSeg observations of one pose per warp, plus that order's copies of the inputs
  std::vector<int> eval_items, eval_item_ptr(h->np + 1, 0), pm_lm_off(nobs_local), pm_cam(nobs_local);
  std::vector<double> pm_pixel(2 * (size_t)nobs_local);
  for (int a = 0; a < h->np; ++a) {
    for (int b = pose_ptr[a]; b < pose_ptr[a + 1]; b += kEvalSeg) { eval_items.push_back(a); eval_items.push_back(b); eval_items.push_back(std::min(b + kEvalSeg, pose_ptr[a + 1])); eval_items.push_back(0); }
    eval_item_ptr[a + 1] = (int)(eval_items.size() / 4);
  }
  h->obs_pm_h.assign(nobs_local, 0);
  ParallelFor(nobs_local, 200000, [&](int i0, int i1) {
    for (int it = i0; it < i1; ++it) {
      const int o = pose_obs[it];
      pm_lm_off[it] = lm_off[s_lm[o]]; pm_cam[it] = s_cam[o]; pm_pixel[2 * (size_t)it] = s_pixel[2 * (size_t)o]; pm_pixel[2 * (size_t)it + 1] = s_pixel[2 * (size_t)o + 1];
      h->obs_pm_h[o] = it;
    }
  });
  h->n_eval_items = (int)(eval_items.size() / 4); h->nb_eval = h->n_eval_items;
  const double avg_track = ne_g ? (double)nobs / ne_g : 0.0;
  // lanes per landmark: throughput-bound shards do best with about three observations per lane (landmark 0.179 -> 0.174 ms
  // on config 4), latency-bound windows with one or two (C1 solve 1.22 vs 1.34 ms)
  if (h->ne >= 20000) h->G = avg_track <= 12.5 ? 4 : avg_track <= 25.0 ? 8 : avg_track <= 50.0 ? 16 : 32;
  else h->G = avg_track <= 5.0 ? 4 : avg_track <= 12.0 ? 8 : avg_track <= 24.0 ? 16 : 32;
  if (const char* ev = std::getenv("BSNLS_LM_LANES")) { const int g = std::atoi(ev); if (g == 4 || g == 8 || g == 16 || g == 32) h->G = g; }   // tuning knob: lanes per landmark

  phase("observation tables");
  // ---- chunks: consecutive landmarks whose poses fit kChunkSlots slots are accumulated on chip by one CTA ----
  std::vector<unsigned short> obs_slot(nobs_local, 0xFFFF);
  std::vector<int> chunk_begin, chunk_end, chunk_nslots, chunk_pose;      // landmark range [begin, end) of each chunk
  std::vector<int> chunk_stage_ptr, stage_lm;                              // stages of a chunk: <= 16 landmarks and <= kFuseMaxObs observations
  std::vector<long long> chunk_tile_off;
  long long tile_size = 0;
  std::vector<char> wide(h->ne, 0);
  {
    std::vector<int> stamp(h->np, -1), cur_poses;
    // chunk size: about 1400 chunks (~5 waves of 2 CTAs x 148 SMs) balances the tail of the last wave against the
    // partial-tile traffic of many small chunks; BSNLS_CHUNK_OBS overrides (tuning knob)
    int chunk_max_obs = std::min(kChunkMaxObs, std::max(512, nobs_local / 1400));
    if (const char* ev = std::getenv("BSNLS_CHUNK_OBS")) chunk_max_obs = std::min(kChunkMaxObs, std::max(64, std::atoi(ev)));
    int c_begin = 0, c_obs = 0, c_kmax = 0;
    auto close_chunk = [&](int end) {
      if (end == c_begin) return;
      if (cur_poses.empty()) { c_begin = end; c_obs = 0; c_kmax = 0; return; }   // only wide landmarks: nothing to accumulate on chip
      std::sort(cur_poses.begin(), cur_poses.end());           // slots in increasing pose id: slot A < B  <=>  pose a < b
      for (size_t q = 0; q < cur_poses.size(); ++q) stamp[cur_poses[q]] = (int)q;
      chunk_stage_ptr.push_back((int)stage_lm.size());
      int st_begin = c_begin, st_obs = 0;
      for (int e = c_begin; e < end; ++e) {
        const int k_e = lm_ptr_local[e + 1] - lm_ptr_local[e];
        if (e > st_begin && (e - st_begin >= kSyrkStageLm || st_obs + k_e > kFuseMaxObs)) { stage_lm.push_back(st_begin); st_begin = e; st_obs = 0; }
        st_obs += k_e;
        if (!wide[e]) for (int k = lm_ptr_local[e]; k < lm_ptr_local[e + 1]; ++k) obs_slot[k] = (unsigned short)(stamp[s_pose[k]] | ((e - st_begin) << 8));
      }
      stage_lm.push_back(st_begin); stage_lm.push_back(end);
      chunk_begin.push_back(c_begin); chunk_end.push_back(end); chunk_nslots.push_back((int)cur_poses.size());
      for (int q = 0; q < kChunkSlots; ++q) chunk_pose.push_back(q < (int)cur_poses.size() ? cur_poses[q] : -1);
      chunk_tile_off.push_back(tile_size);
      { const long long nt2 = (6 * (long long)cur_poses.size() + 8) >> 3; tile_size += nt2 * (nt2 + 1) / 2 * 64; }
      for (int pz : cur_poses) stamp[pz] = -1;
      cur_poses.clear(); c_begin = end; c_obs = 0; c_kmax = 0;
    };
    for (int e = 0; e < h->ne; ++e) {
      const int k0 = lm_ptr_local[e], k1 = lm_ptr_local[e + 1];
      // wide landmark: more poses than slots, or seen twice by one pose -> handled through the item list
      // small shards: the on-chip chunk path needs a few hundred chunks to fill the GPU; below that every landmark goes
      // through the item list (warp-per-block pair_kernel)
      if (e - c_begin >= kChunkMaxLm) close_chunk(e);
      bool is_wide = k1 - k0 > kChunkSlots || nobs_local < kChunkMinObs;
      for (int k = k0; k < k1 && !is_wide; ++k) for (int k2 = k0; k2 < k; ++k2) if (s_pose[k2] == s_pose[k]) { is_wide = true; break; }
      if (is_wide) { wide[e] = 1; continue; }
      int fresh = 0;
      for (int k = k0; k < k1; ++k) if (stamp[s_pose[k]] == -1) ++fresh;
      const int limit = std::min(kChunkSlots, std::max(c_kmax, k1 - k0) + kChunkSlack);
      if ((int)cur_poses.size() + fresh > limit || c_obs + (k1 - k0) > chunk_max_obs) { close_chunk(e); }
      c_kmax = std::max(c_kmax, k1 - k0);
      for (int k = k0; k < k1; ++k) if (stamp[s_pose[k]] == -1) { stamp[s_pose[k]] = -3; cur_poses.push_back(s_pose[k]); }
      c_obs += k1 - k0;
    }
    // note: stamp == -3 marks "in the current chunk" while it is being built
    close_chunk(h->ne);
  }
  h->nchunks = (int)chunk_nslots.size();
  chunk_stage_ptr.push_back((int)stage_lm.size());
  h->chunk_max_slots = 1; for (int v : chunk_nslots) h->chunk_max_slots = std::max(h->chunk_max_slots, v);
  // chunk-major fused pass (chunk_fused_kernel), opt-in with BSNLS_FUSED=1: it moves 2.2x fewer DRAM bytes than the
  // three-kernel path but is slower on B200 (its phases serialise inside a CTA). Every landmark of the
  // shard must go through a chunk.
  { bool any_wide = false; for (char w : wide) any_wide = any_wide || w;
    h->fused = h->nchunks > 0 && !any_wide && std::getenv("BSNLS_FUSED") != nullptr; }
  std::vector<int> pc_ptr(h->np + 1, 0), pc_idx;   // per pose: its (chunk, slot) partial sums
  if (h->fused) {
    for (int c = 0; c < h->nchunks; ++c) for (int q = 0; q < chunk_nslots[c]; ++q) pc_ptr[chunk_pose[(size_t)c * kChunkSlots + q] + 1]++;
    for (int a = 0; a < h->np; ++a) pc_ptr[a + 1] += pc_ptr[a];
    pc_idx.resize(pc_ptr[h->np]);
    std::vector<int> cur(pc_ptr.begin(), pc_ptr.end() - 1);
    for (int c = 0; c < h->nchunks; ++c) for (int q = 0; q < chunk_nslots[c]; ++q) pc_idx[cur[chunk_pose[(size_t)c * kChunkSlots + q]]++] = c * kChunkSlots + q;
    for (int it = 0; it < nobs_local; ++it) h->obs_pm_h[it] = it;   // Jacobian records live at their landmark-major position
    h->nb_eval = h->nchunks;                                         // one cost partial per chunk
  }
  if (std::getenv("BSNLS_DEBUG")) {
    long long slots = 0, nwide = 0; for (int v : chunk_nslots) slots += v; for (char w : wide) nwide += w;
    std::fprintf(stderr, "[bsnls] rank %d: %d landmarks, %d obs, %d poses, %d pairs, %d chunks (avg %.1f slots, %.1f landmarks), %lld wide landmarks, partial buffer %.1f MB\n",
                 h->rank, h->ne, nobs_local, h->np, h->npairs, h->nchunks, h->nchunks ? (double)slots / h->nchunks : 0.0, h->nchunks ? (double)h->ne / h->nchunks : 0.0, nwide, tile_size * 8e-6);
  }
  phase("chunks");
  // ---- per pair: partial blocks produced by the chunks, and items of the wide landmarks ----
  std::vector<int> pair_ptr(h->npairs + 1, 0), item_i, item_j, pp_ptr(h->npairs + 1, 0), local_pairs;
  std::vector<long long> pp_toff; std::vector<int> pp_info;
  {
    std::vector<int> cnt(h->npairs, 0), pcnt(h->npairs, 0);
    for (int e = 0; e < h->ne; ++e) {
      if (!wide[e]) continue;
      for (int i = lm_ptr_local[e]; i < lm_ptr_local[e + 1]; ++i)
        for (int j = i; j < lm_ptr_local[e + 1]; ++j) cnt[pair_id(s_pose[i], s_pose[j])] += (s_pose[i] == s_pose[j] && i != j) ? 2 : 1;
    }
    for (int c = 0; c < h->nchunks; ++c) {
      const int ns = chunk_nslots[c];
      for (int A = 0; A < ns; ++A) for (int B = A; B < ns; ++B) {
        // two poses of a chunk need not share a landmark: such a slot pair is not a block of S (its partial block stays zero)
        const int id = pair_id(chunk_pose[(size_t)c * kChunkSlots + A], chunk_pose[(size_t)c * kChunkSlots + B]);
        if (id >= 0) pcnt[id]++;
      }
    }
    for (int k = 0; k < h->npairs; ++k) { pair_ptr[k + 1] = pair_ptr[k] + cnt[k]; pp_ptr[k + 1] = pp_ptr[k] + pcnt[k]; }
    const int64_t nitems = pair_ptr[h->npairs];
    item_i.resize(nitems); item_j.resize(nitems); pp_toff.resize(pp_ptr[h->npairs]); pp_info.resize(pp_ptr[h->npairs]);
    std::vector<int> cur(pair_ptr.begin(), pair_ptr.end() - 1), pcur(pp_ptr.begin(), pp_ptr.end() - 1);
    for (int e = 0; e < h->ne; ++e) {
      if (!wide[e]) continue;
      for (int i = lm_ptr_local[e]; i < lm_ptr_local[e + 1]; ++i)
        for (int j = i; j < lm_ptr_local[e + 1]; ++j) {
          const int a = s_pose[i], b = s_pose[j];
          const int id = pair_id(a, b);
          if (a < b || i == j) { item_i[cur[id]] = i; item_j[cur[id]] = j; cur[id]++; }
          else if (a > b) { item_i[cur[id]] = j; item_j[cur[id]] = i; cur[id]++; }
          else { item_i[cur[id]] = i; item_j[cur[id]] = j; cur[id]++; item_i[cur[id]] = j; item_j[cur[id]] = i; cur[id]++; }
        }
    }
    for (int c = 0; c < h->nchunks; ++c) {
      const int ns = chunk_nslots[c];
      int q = 0;
      for (int A = 0; A < ns; ++A) for (int B = A; B < ns; ++B, ++q) {
        const int id = pair_id(chunk_pose[(size_t)c * kChunkSlots + A], chunk_pose[(size_t)c * kChunkSlots + B]);
        if (id < 0) continue;
        pp_toff[pcur[id]] = chunk_tile_off[c]; pp_info[pcur[id]] = A | (B << 8) | (ns << 16);
        pcur[id]++;
      }
    }
    h->nitems = nitems;
    local_pairs.clear();
    for (int k = 0; k < h->npairs; ++k) if (pair_ptr[k + 1] > pair_ptr[k] || pp_ptr[k + 1] > pp_ptr[k]) local_pairs.push_back(k);
    h->n_local_pairs = (int)local_pairs.size();
    if (std::getenv("BSNLS_DEBUG")) {
      int mx = 0, mxc = 0; long long tot = 0; for (int k = 0; k < h->npairs; ++k) { mx = std::max(mx, pp_ptr[k + 1] - pp_ptr[k]); tot += pp_ptr[k + 1] - pp_ptr[k]; }
      for (int c = 0; c < h->nchunks; ++c) mxc = std::max(mxc, chunk_end[c] - chunk_begin[c]);
      std::fprintf(stderr, "[bsnls] partial lists: total %lld, max per pair %d, avg %.2f; items %lld; largest chunk %d landmarks\n", tot, mx, h->npairs ? (double)tot / h->npairs : 0.0, (long long)nitems, mxc);
    }
  }

  phase("pair partial lists");
  // ---- generic factors (type-major; factors whose blocks are all constant go last: fixed cost) ----
  std::vector<int> gf_type, gf_nblk, gf_nres, gf_blocks, gf_loss_type, gf_joff, gf_roff;
  std::vector<long long> gf_coff; std::vector<double> gf_consts, gf_loss_a;
  struct GfRef { const bsnls_factor_group* g; int i; bool removed; };
  std::vector<GfRef> refs;
  for (const bsnls_factor_group* G : gfg) {
    const int nbk = kNumBlocks[G->type];
    for (int i = 0; i < G->count; ++i) {
      bool any = false;
      for (int b = 0; b < UsedBlocks(*G, i); ++b) {
        const int blk = G->blocks[(int64_t)i * nbk + b];
        any = any || !pr->block_constant[blk];
        if (epos[blk] >= 0) return h->fail(BSNLS_ERR_UNSUPPORTED, "a non-reprojection factor references an elimination-group-0 block");
      }
      refs.push_back({G, i, !any});
    }
  }
  std::stable_sort(refs.begin(), refs.end(), [](const GfRef& a, const GfRef& b) { return a.removed < b.removed; });
  h->ngf = (int)refs.size(); h->ngf_active = 0;
  std::vector<std::vector<int>> gf_id(BSNLS_NUM_FACTOR_TYPES);  // [type][index in group] -> gf id
  for (const bsnls_factor_group* G : gfg) gf_id[G->type].assign(G->count, -1);
  int64_t joff = 0, roff = 0;
  for (int f = 0; f < h->ngf; ++f) {
    const bsnls_factor_group* G = refs[f].g; const int i = refs[f].i, slots = kNumBlocks[G->type], nbk = UsedBlocks(*G, i), nr = kNumRes[G->type], stride = kStride[G->type];
    if (!refs[f].removed) h->ngf_active++;
    gf_id[G->type][i] = f;
    gf_type.push_back(G->type); gf_nblk.push_back(nbk); gf_nres.push_back(nr);
    for (int b = 0; b < kGfMaxBlocks; ++b) gf_blocks.push_back(b < nbk ? G->blocks[(int64_t)i * slots + b] : -1);
    gf_coff.push_back((long long)gf_consts.size());
    gf_consts.insert(gf_consts.end(), G->consts + (int64_t)i * stride, G->consts + (int64_t)(i + 1) * stride);
    gf_loss_type.push_back(G->loss_type ? G->loss_type[i] : 0); gf_loss_a.push_back(G->loss_param ? G->loss_param[i] : 0.0);
    gf_joff.push_back((int)joff); gf_roff.push_back((int)roff);
    joff += (int64_t)nr * 3 * nbk; roff += nr;
  }
  h->gfJ_size = joff; h->gfR_size = roff; h->gf_roff_h = gf_roff; h->gf_nres_h = gf_nres; h->gf_joff_h = gf_joff; h->gf_nblk_h = gf_nblk;
  // contributions per f-block and per S destination
  std::vector<int> fg_ptr(h->nf + 1, 0), fg_fac, fg_lb;
  {
    std::vector<std::vector<std::pair<int, int>>> per(h->nf);
    for (int f = 0; f < h->ngf_active; ++f)
      for (int b = 0; b < gf_nblk[f]; ++b) { const int blk = gf_blocks[f * kGfMaxBlocks + b]; if (fpos[blk] >= 0) per[fpos[blk]].push_back({f, b}); }
    for (int k = 0; k < h->nf; ++k) { fg_ptr[k + 1] = fg_ptr[k] + (int)per[k].size(); for (auto& pb : per[k]) { fg_fac.push_back(pb.first); fg_lb.push_back(pb.second); } }
  }
  std::vector<int> gd_ptr, gd_fac, gd_bi, gd_bj; h->gd_row.clear(); h->gd_col.clear();
  {
    struct C { uint64_t key; int f, bi, bj; };
    std::vector<C> cs;
    for (int f = 0; f < h->ngf_active; ++f)
      for (int bi = 0; bi < gf_nblk[f]; ++bi)
        for (int bj = 0; bj < gf_nblk[f]; ++bj) {
          const int R = fpos[gf_blocks[f * kGfMaxBlocks + bi]], Cc = fpos[gf_blocks[f * kGfMaxBlocks + bj]];
          if (R < 0 || Cc < 0 || R < Cc) continue;
          if (R == Cc && bi != bj) continue;
          cs.push_back({(uint64_t)R * (uint64_t)h->nf + (uint64_t)Cc, f, bi, bj});
        }
    std::stable_sort(cs.begin(), cs.end(), [](const C& a, const C& b) { return a.key < b.key; });
    gd_ptr.push_back(0);
    for (size_t k = 0; k < cs.size(); ++k) {
      if (k == 0 || cs[k].key != cs[k - 1].key) { if (k) gd_ptr.push_back((int)k); h->gd_row.push_back((int)(cs[k].key / h->nf)); h->gd_col.push_back((int)(cs[k].key % h->nf)); }
      gd_fac.push_back(cs[k].f); gd_bi.push_back(cs[k].bi); gd_bj.push_back(cs[k].bj);
    }
    gd_ptr.push_back((int)cs.size());
    h->ngdest = (int)h->gd_row.size();
    if (cs.empty()) gd_ptr.assign(1, 0);
  }
  std::vector<unsigned char> is_pose(nf_dim, 0);
  for (int a = 0; a < h->np; ++a) {
    if (pose_q_tan[a] >= 0) for (int k = 0; k < 3; ++k) is_pose[pose_q_tan[a] - ne_dim + k] = 1;
    if (pose_t_tan[a] >= 0) for (int k = 0; k < 3; ++k) is_pose[pose_t_tan[a] - ne_dim + k] = 1;
  }
  // ---- block-sparse pattern of the reduced system (3x3 blocks over f-blocks, full symmetric) ----
  std::vector<int> low_slot, low_slot_t;
  std::vector<int> diag_slot(h->nf, 0), pair_slot((size_t)h->npairs * 8, -1), gd_slot((size_t)h->ngdest * 2, -1);
  h->bsr_rowptr.assign(h->nf + 1, 0); h->bsr_col.clear(); h->nnzb = 0;
  {
    std::vector<uint64_t> keys; keys.reserve((size_t)h->npairs * 8 + (size_t)h->ngdest * 2 + h->nf);
    auto add = [&](int r, int c) { if (r >= 0 && c >= 0) { keys.push_back((uint64_t)r * h->nf + c); keys.push_back((uint64_t)c * h->nf + r); } };
    auto fq = [&](int a) { return fpos[h->pose_q_blk[a]]; }; auto ft = [&](int a) { return fpos[h->pose_t_blk[a]]; };
    for (int k = 0; k < h->nf; ++k) add(k, k);
    for (int p2 = 0; p2 < h->npairs; ++p2) { const int a = h->pair_a[p2], b = h->pair_b[p2]; add(fq(a), fq(b)); add(fq(a), ft(b)); add(ft(a), fq(b)); add(ft(a), ft(b)); }
    for (int d = 0; d < h->ngdest; ++d) add(h->gd_row[d], h->gd_col[d]);
    std::sort(keys.begin(), keys.end()); keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
    h->nnzb = (int64_t)keys.size();
    if (h->nnzb * 9 > 0x7fffffffLL) return h->fail(BSNLS_ERR_UNSUPPORTED, "reduced system has too many blocks");
    h->bsr_col.resize(keys.size());
    for (size_t k = 0; k < keys.size(); ++k) { const int r = (int)(keys[k] / h->nf), c = (int)(keys[k] % h->nf); h->bsr_rowptr[r + 1]++; h->bsr_col[k] = c; }
    for (int r = 0; r < h->nf; ++r) h->bsr_rowptr[r + 1] += h->bsr_rowptr[r];
    auto slot = [&](int r, int c) -> int {
      if (r < 0 || c < 0) return -1;
      const int* b0 = h->bsr_col.data() + h->bsr_rowptr[r]; const int* b1 = h->bsr_col.data() + h->bsr_rowptr[r + 1];
      return (int)(std::lower_bound(b0, b1, c) - h->bsr_col.data());
    };
    for (int k = 0; k < h->nf; ++k) diag_slot[k] = slot(k, k);
    for (int p2 = 0; p2 < h->npairs; ++p2) {
      const int a = h->pair_a[p2], b = h->pair_b[p2];
      const int ra[2] = {fq(a), ft(a)}, cb[2] = {fq(b), ft(b)};
      for (int sa = 0; sa < 2; ++sa) for (int sb = 0; sb < 2; ++sb) {
        pair_slot[(size_t)p2 * 8 + 2 * sa + sb] = slot(ra[sa], cb[sb]);
        pair_slot[(size_t)p2 * 8 + 4 + 2 * sa + sb] = slot(cb[sb], ra[sa]);
      }
    }
    for (int d = 0; d < h->ngdest; ++d) { gd_slot[2 * (size_t)d] = slot(h->gd_row[d], h->gd_col[d]); gd_slot[2 * (size_t)d + 1] = slot(h->gd_col[d], h->gd_row[d]); }
    if (h->world > 1)
      for (int r = 0; r < h->nf; ++r)
        for (int k = h->bsr_rowptr[r]; k < h->bsr_rowptr[r + 1]; ++k) {
          const int c = h->bsr_col[k];
          if (c <= r) { low_slot.push_back(k); low_slot_t.push_back(c < r ? slot(c, r) : -1); }
        }
  }
  // residual order for bsnls_evaluate
  h->res_order.assign(total, {0, 0, 0});
  {
    std::vector<char> seen(total, 0);
    for (int g = 0; g < pr->num_groups; ++g) {
      const bsnls_factor_group& G = pr->groups[g];
      for (int i = 0; i < G.count; ++i) {
        const int ri = G.res_index[i];
        if (ri < 0 || ri >= total || seen[ri]) return h->fail(BSNLS_ERR_INVALID, "res_index is not a permutation of 0..N-1");
        seen[ri] = 1;
        h->res_order[ri] = {G.type, G.type == BSNLS_FACTOR_REPROJECTION ? inv_order[i] : gf_id[G.type][i], 0};
      }
    }
    int row = 0;
    for (auto& r : h->res_order) { r.row = row; row += kNumRes[r.type]; }
    h->num_res_rows = row;
  }

  phase("generic factors + S pattern");
  // ---- upload: one arena, one staged copy (see Arena) ----
  cudaStream_t s = h->stream;
  struct RecGuard { ~RecGuard() { g_rec = nullptr; } } rec_guard;
  if (h->stage_busy) { CK(cudaEventSynchronize(h->ev_stage)); h->stage_busy = false; }   // the previous problem's staged copy
  h->arena.Begin(); g_rec = &h->arena;
  CK(h->arena.StageReserve(std::min<size_t>((size_t)nobs_local * 72 + (size_t)nb * 48 + (size_t)off * 9 + (size_t)h->npairs * 64, size_t(24) << 20) + (size_t(1) << 20)));   // one pinned allocation for the small arrays (arrays above 1 MB bypass the mirror)
  CK(upload(h->d_blk_off, h->blk_off, s)); CK(upload(h->d_blk_size, h->blk_size, s)); CK(upload(h->d_blk_tan, h->blk_tan, s));
  CK(upload(h->d_blk_manifold, h->blk_manifold, s)); CK(upload(h->d_active, h->active_blocks, s));
  CK(upload(h->d_pose_q_off, pose_q_off, s)); CK(upload(h->d_pose_t_off, pose_t_off, s)); CK(upload(h->d_pose_q_tan, pose_q_tan, s));
  CK(upload(h->d_pose_t_tan, pose_t_tan, s)); CK(upload(h->d_pose_ptr, pose_ptr, s)); CK(upload(h->d_pose_obs, pose_obs, s));
  CK(upload(h->d_obs_pose, s_pose, s)); CK(upload(h->d_obs_lm, s_lm, s)); CK(upload(h->d_obs_cam, s_cam, s)); CK(upload(h->d_obs_pixel, s_pixel, s));
  CK(upload(h->d_lm_ptr, lm_ptr_local, s)); CK(upload(h->d_lm_off, lm_off, s));
  CK(upload(h->d_obs_pm, h->obs_pm_h, s)); CK(upload(h->d_eval_items, eval_items, s)); CK(upload(h->d_eval_item_ptr, eval_item_ptr, s)); CK(upload(h->d_pm_lm_off, pm_lm_off, s));
  CK(upload(h->d_pm_cam, pm_cam, s)); CK(upload(h->d_pm_pixel, pm_pixel, s));
  CK(h->d_pose_tab.alloc(sizeof(double) * kPoseTab * (size_t)std::max(1, h->np))); CK(h->d_pose_part.alloc(sizeof(double) * 27 * (size_t)std::max(1, h->fused ? h->nchunks * kChunkSlots : h->n_eval_items)));
  { std::vector<double> cams(pr->cameras, pr->cameras + (size_t)pr->num_cameras * BSNLS_CAMERA_STRIDE); CK(upload(h->d_cams, cams, s)); }
  CK(upload(h->d_gf_type, gf_type, s)); CK(upload(h->d_gf_nblk, gf_nblk, s)); CK(upload(h->d_gf_nres, gf_nres, s)); CK(upload(h->d_gf_blocks, gf_blocks, s));
  CK(upload(h->d_gf_coff, gf_coff, s)); CK(upload(h->d_gf_consts, gf_consts, s)); CK(upload(h->d_gf_loss_type, gf_loss_type, s));
  CK(upload(h->d_gf_loss_a, gf_loss_a, s)); CK(upload(h->d_gf_joff, gf_joff, s)); CK(upload(h->d_gf_roff, gf_roff, s));
  CK(upload(h->d_pair_a, h->pair_a, s)); CK(upload(h->d_pair_b, h->pair_b, s)); CK(upload(h->d_pair_ptr, pair_ptr, s));
  CK(upload(h->d_item_i, item_i, s)); CK(upload(h->d_item_j, item_j, s));
  CK(upload(h->d_obs_slot, obs_slot, s)); CK(upload(h->d_chunk_begin, chunk_begin, s)); CK(upload(h->d_chunk_end, chunk_end, s)); CK(upload(h->d_chunk_nslots, chunk_nslots, s));
  CK(upload(h->d_pp_ptr, pp_ptr, s)); CK(upload(h->d_local_pairs, local_pairs, s));
  CK(upload(h->d_chunk_pose, chunk_pose, s)); CK(upload(h->d_chunk_stage_ptr, chunk_stage_ptr, s)); CK(upload(h->d_stage_lm, stage_lm, s));
  CK(upload(h->d_pc_ptr, pc_ptr, s)); CK(upload(h->d_pc_idx, pc_idx, s));
  { std::vector<int> plan((kSyrkMaxNt + 1) * 8 * kSyrkMaxStrips); syrk_build_plan(plan.data()); CK(upload(h->d_syrk_plan, plan, s)); }
  CK(upload(h->d_chunk_tile_off, chunk_tile_off, s)); CK(upload(h->d_pp_toff, pp_toff, s)); CK(upload(h->d_pp_info, pp_info, s));
  CK(h->d_part.alloc(sizeof(double) * (size_t)tile_size));
  CK(upload(h->d_fg_ptr, fg_ptr, s)); CK(upload(h->d_fg_fac, fg_fac, s)); CK(upload(h->d_fg_lb, fg_lb, s));
  CK(upload(h->d_gd_row, h->gd_row, s)); CK(upload(h->d_gd_col, h->gd_col, s)); CK(upload(h->d_gd_ptr, gd_ptr, s));
  h->nlow = (long long)low_slot.size();
  if (h->world > 1 && h->sparse) {
    CK(upload(h->d_low_slot, low_slot, s)); CK(upload(h->d_low_slot_t, low_slot_t, s));
    const size_t xn = 9 * (size_t)h->nlow + 3 * (size_t)nf_dim + RED_NUM + h->world + 1;
    CK(h->d_xsend.alloc(sizeof(double) * xn)); CK(h->d_xrecv.alloc(sizeof(double) * xn));
  }
  CK(upload(h->d_gd_fac, gd_fac, s)); CK(upload(h->d_gd_bi, gd_bi, s)); CK(upload(h->d_gd_bj, gd_bj, s)); CK(upload(h->d_is_pose, is_pose, s));
  CK(upload(h->d_bsr_rowptr, h->bsr_rowptr, s)); CK(upload(h->d_bsr_col, h->bsr_col, s)); CK(upload(h->d_diag_slot, diag_slot, s));
  CK(upload(h->d_pair_slot, pair_slot, s)); CK(upload(h->d_gd_slot, gd_slot, s));
  std::vector<unsigned char> owned;   // function scope: large uploads are copied at Commit
  { // ownership of parameter entries for the final gather of the sharded solve
    owned.assign((size_t)off, h->rank == 0 ? 1 : 0);
    for (int b = 0; b < nb; ++b) if (epos[b] >= 0) { const unsigned char own = is_mine[epos[b]] ? 1 : 0; for (int k = 0; k < h->blk_size[b]; ++k) owned[h->blk_off[b] + k] = own; }
    CK(upload(h->d_owned, owned, s));
  }
  phase("staging of uploads");
  // work buffers
  const int ntan = ne_dim + nf_dim;
  // ---- tile-sparse Cholesky: symbolic factorisation at tile granularity (structure only; fixed per problem) ----
  // A tile holds kTileBlocks = 21 f-blocks (63 unknowns, the 64th row/column is an identity pad), so tiles never cut a block.
  std::vector<long long> tc_tile_off; std::vector<int> tc_col_rows; long long tc_tiles = 0;
  std::vector<int> tc_pan_k, tc_pan_x, tc_upd_ptr, tc_upd_fac, tc_fw_ptr, tc_fw_row, tc_fw_tail, tc_lev_cols;
  std::vector<long long> tc_upd_dst, tc_upd_a, tc_upd_b, tc_fw_src; std::vector<int> tc_fw_k;
  std::vector<int> tc_bw_ptr, tc_bw_col, tc_bw_tail, tc_bw_i, tc_roots, tc_slot_blk, tc_row_ptr, tc_row_tile; std::vector<long long> tc_bw_src;
  h->tc_col_ptr.clear(); h->tc_levels.clear();
  if (h->tiled) {
    const int nT = h->tc_nT;
    std::vector<std::vector<int>> col(nT);     // non-zero tiles strictly below the diagonal, per tile column
    for (int r = 0; r < h->nf; ++r)
      for (int k = h->bsr_rowptr[r]; k < h->bsr_rowptr[r + 1]; ++k) {
        const int c = h->bsr_col[k];
        if (c < r && h->tc_blk_slot[r] / kTileBlocks > h->tc_blk_slot[c] / kTileBlocks) col[h->tc_blk_slot[c] / kTileBlocks].push_back(h->tc_blk_slot[r] / kTileBlocks);
      }
    std::vector<int> level(nT, 0);
    for (int k = 0; k < nT; ++k) {
      std::sort(col[k].begin(), col[k].end()); col[k].erase(std::unique(col[k].begin(), col[k].end()), col[k].end());
      // fill: eliminating tile column k couples all of its rows; it is enough to hand the rows below the first one to
      // that first row's column (elimination-tree parent): later columns propagate them further
      if (!col[k].empty()) { const int par = col[k][0]; col[par].insert(col[par].end(), col[k].begin() + 1, col[k].end()); level[par] = std::max(level[par], level[k] + 1); }
    }
    h->tc_col_ptr.assign(nT + 1, 0);
    for (int k = 0; k < nT; ++k) { h->tc_col_ptr[k + 1] = h->tc_col_ptr[k] + (int)col[k].size(); tc_tiles += 1 + (long long)col[k].size(); }
    const double tile_gb = (double)tc_tiles * kTileDoubles * 8 / 1e9;
    const int nlev = 1 + *std::max_element(level.begin(), level.end());
    if (std::getenv("BSNLS_DEBUG")) std::fprintf(stderr, "[bsnls] tile Cholesky: n %d, %d tile columns in %d levels, %lld tiles (%.3f GB), dense lower would be %lld\n", nf_dim, nT, nlev, tc_tiles, tile_gb, (long long)nT * (nT + 1) / 2);
    if (tile_gb > 24.0 || (long long)nT * nT > (1LL << 28)) {
      if (h->opt.linear_solver == BSNLS_LINEAR_DENSE_SCHUR) return h->fail(BSNLS_ERR_UNSUPPORTED, "tile factor of the reduced system needs %.1f GB; use the iterative solver", tile_gb);
      h->tiled = false; h->pcg = true;     // AUTO: fall back to conjugate gradients
    } else {
      tc_tile_off.assign((size_t)nT * nT, -1);
      long long t = 0;
      for (int k = 0; k < nT; ++k) {
        tc_tile_off[(size_t)k * nT + k] = (t++) * kTileDoubles;
        for (int i : col[k]) { tc_tile_off[(size_t)i * nT + k] = (t++) * kTileDoubles; tc_col_rows.push_back(i); }
      }
      { // rows of the factor: the tiles (k, j), j < k, for the forward substitution inside the panel kernels
        std::vector<std::vector<int>> rows(nT);
        for (int j = 0; j < nT; ++j) for (int i : col[j]) rows[i].push_back((int)(tc_tile_off[(size_t)i * nT + j] / kTileDoubles));
        tc_row_ptr.assign(nT + 1, 0);
        for (int k = 0; k < nT; ++k) { tc_row_ptr[k + 1] = tc_row_ptr[k] + (int)rows[k].size(); tc_row_tile.insert(tc_row_tile.end(), rows[k].begin(), rows[k].end()); }
        if (tc_row_tile.empty()) tc_row_tile.push_back(0);
      }
      // level schedule: tile columns of one level of the elimination tree are independent. Per level: the panel CTAs
      // (column, tile below the diagonal), the destination tiles of the updates with their source columns, and for
      // the forward substitution the destination rows with their source tiles.
      std::vector<std::vector<int>> by_level(nlev);
      for (int k = 0; k < nT; ++k) by_level[level[k]].push_back(k);
      tc_upd_ptr.push_back(0); tc_fw_ptr.push_back(0); tc_bw_ptr.push_back(0);
      for (int k = 0; k < nT; ++k) if (col[k].empty()) tc_roots.push_back(k);   // backward substitution starts from the columns without rows
      // backward: per level (of the SOURCE rows) the destination columns that have rows on it
      std::vector<std::vector<std::pair<int, int>>> bw_by_level(nlev);   // (dest column, source row)
      for (int k = 0; k < nT; ++k) for (int i : col[k]) bw_by_level[level[i]].push_back({k, i});
      for (int lv = 0; lv < nlev; ++lv) {
        bsnls_handle::TcLevel L{};
        L.col_begin = (int)tc_lev_cols.size(); L.pan_begin = (int)tc_pan_k.size(); L.upd_begin = (int)tc_upd_ptr.size() - 1; L.fw_begin = (int)tc_fw_ptr.size() - 1;
        struct U { long long key; int k; }; std::vector<U> us;   // key: destinations that are factored afterwards first, then (row, column)
        const long long kPlain = 1LL << 62; struct F { int row, k; }; std::vector<F> fs;
        for (int k : by_level[lv]) {
          tc_lev_cols.push_back(k);
          const int a = (int)col[k].size();
          if (a == 0) { tc_pan_k.push_back(k); tc_pan_x.push_back(-1); }
          for (int x = 0; x < a; ++x) { tc_pan_k.push_back(k); tc_pan_x.push_back(h->tc_col_ptr[k] + x); fs.push_back({col[k][x], k}); for (int y2 = 0; y2 <= x; ++y2) us.push_back({(long long)col[k][x] * nT + col[k][y2] + ((y2 == x && level[col[k][x]] == lv + 1) ? 0 : kPlain), k}); }
        }
        std::stable_sort(us.begin(), us.end(), [](const U& a2, const U& b2) { return a2.key < b2.key; });
        for (size_t q = 0; q < us.size(); ++q) {
          const bool fac = us[q].key < kPlain; const long long key = fac ? us[q].key : us[q].key - kPlain;
          const int i = (int)(key / nT), j = (int)(key % nT), k = us[q].k;
          if (q == 0 || us[q].key != us[q - 1].key) { if (q) tc_upd_ptr.push_back((int)tc_upd_a.size()); tc_upd_dst.push_back(tc_tile_off[(size_t)i * nT + j]); tc_upd_fac.push_back(fac ? i : -1); }
          tc_upd_a.push_back(tc_tile_off[(size_t)i * nT + k]); tc_upd_b.push_back(tc_tile_off[(size_t)j * nT + k]);
        }
        if (!us.empty()) tc_upd_ptr.push_back((int)tc_upd_a.size());
        std::stable_sort(fs.begin(), fs.end(), [](const F& a2, const F& b2) { return a2.row < b2.row; });
        for (size_t q = 0; q < fs.size(); ++q) {
          if (q == 0 || fs[q].row != fs[q - 1].row) { if (q) tc_fw_ptr.push_back((int)tc_fw_src.size()); tc_fw_row.push_back(fs[q].row); tc_fw_tail.push_back(level[fs[q].row] == lv + 1 ? 1 : 0); }
          tc_fw_src.push_back(tc_tile_off[(size_t)fs[q].row * nT + fs[q].k]); tc_fw_k.push_back(fs[q].k);
        }
        if (!fs.empty()) tc_fw_ptr.push_back((int)tc_fw_src.size());
        L.ncols = (int)tc_lev_cols.size() - L.col_begin; L.npan = (int)tc_pan_k.size() - L.pan_begin;
        L.npan_tiles = 0; for (int q = L.pan_begin; q < L.pan_begin + L.npan; ++q) L.npan_tiles += tc_pan_x[q] >= 0 ? 1 : 0;
        L.nupd = (int)tc_upd_ptr.size() - 1 - L.upd_begin; L.nfw = (int)tc_fw_ptr.size() - 1 - L.fw_begin;
        L.bw_begin = (int)tc_bw_ptr.size() - 1;
        auto& bw = bw_by_level[lv];       // already grouped by destination column (k ascending), rows ascending
        for (size_t q = 0; q < bw.size(); ++q) {
          const int k = bw[q].first, i = bw[q].second;
          if (q == 0 || bw[q - 1].first != k) { if (q) tc_bw_ptr.push_back((int)tc_bw_src.size()); tc_bw_col.push_back(k); tc_bw_tail.push_back(level[col[k][0]] == lv ? 1 : 0); }
          tc_bw_src.push_back(tc_tile_off[(size_t)i * nT + k]); tc_bw_i.push_back(i);
        }
        if (!bw.empty()) tc_bw_ptr.push_back((int)tc_bw_src.size());
        L.nbw = (int)tc_bw_ptr.size() - 1 - L.bw_begin;
        h->tc_levels.push_back(L);
        if (std::getenv("BSNLS_DEBUG")) std::fprintf(stderr, "[bsnls]   level %d: %d columns, %d panel CTAs, %d update tiles (%d products)\n", lv, L.ncols, L.npan, L.nupd, (int)tc_upd_a.size() - tc_upd_ptr[L.upd_begin]);
      }
    }
  }
  phase("tile symbolic factorisation");
  h->nb_obs = h->nobs ? BlocksFor(h->nobs, 256) : 0; h->nb_lm = h->ne ? BlocksFor(h->ne * h->G, 256) : 0; h->nb_act = BlocksFor((int)h->active_blocks.size(), 256);
  { // parameter values travel with the staged structure copy
    cudaError_t e = g_rec->StageReserve(g_rec->up_used + Arena::Align(sizeof(double) * off)); if (e != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "staging: %s", cudaGetErrorString(e));
    DevMem& m = h->d_x; if (m.p && !m.view) cudaFree(m.p);
    m.p = nullptr; m.view = true; m.in_work = false; m.bytes = sizeof(double) * off; m.off = g_rec->up_used;
    std::memcpy(g_rec->stage + m.off, pr->block_values, sizeof(double) * off);
    g_rec->up_used += Arena::Align(sizeof(double) * off); g_rec->members.push_back(&m);
  }
  CK(h->d_xc.alloc(sizeof(double) * off)); CK(h->d_x0.alloc(sizeof(double) * off));
  CK(h->d_Jrec.alloc(sizeof(double) * kJRec * (size_t)h->nobs)); CK(h->d_Vrec.alloc(h->fused ? 8 : sizeof(double) * kVRec * (size_t)h->nobs));   // the fused pass keeps V on chip
  CK(h->d_minv.alloc(sizeof(double) * 6 * (size_t)h->ne)); CK(h->d_lmz.alloc(sizeof(double) * 3 * (size_t)h->ne)); CK(h->d_grad.alloc(sizeof(double) * ntan)); CK(h->d_colsq.alloc(sizeof(double) * ntan));
  CK(h->d_zero.alloc(sizeof(double) * (2 * (size_t)nf_dim + ntan)));   // gradL | colsqL | y: start at zero, cleared by one memset
  if (h->world > 1) { CK(h->d_gradS.alloc(sizeof(double) * nf_dim)); CK(h->d_colsqS.alloc(sizeof(double) * nf_dim)); }
  CK(h->d_scale.alloc(sizeof(double) * ntan)); CK(h->d_yf.alloc(sizeof(double) * nf_dim));
  CK(h->d_rhs.alloc(sizeof(double) * nf_dim)); CK(h->d_S.alloc(sizeof(double) * (h->sparse ? (size_t)h->nnzb * 9 : (size_t)nf_dim * nf_dim)));
  if (h->tiled) {
    const int nT = h->tc_nT;
    tc_slot_blk.assign((size_t)nT * kTileBlocks, -1);
    for (int r = 0; r < h->nf; ++r) tc_slot_blk[h->tc_blk_slot[r]] = r;
    CK(upload(h->d_tc_blk_slot, h->tc_blk_slot, s)); CK(upload(h->d_tc_slot_blk, tc_slot_blk, s));
    CK(upload(h->d_tc_row_ptr, tc_row_ptr, s)); CK(upload(h->d_tc_row_tile, tc_row_tile, s));
    CK(h->d_tc_braw.alloc(sizeof(double) * kTS * (size_t)nT)); CK(h->d_tc_P.alloc(sizeof(double) * kTS * (size_t)tc_tiles));
    if (tc_col_rows.empty()) tc_col_rows.push_back(0);
    if (tc_upd_dst.empty()) { tc_upd_dst.push_back(0); tc_upd_fac.push_back(-1); tc_upd_a.push_back(0); tc_upd_b.push_back(0); }
    if (tc_fw_row.empty()) { tc_fw_row.push_back(0); tc_fw_tail.push_back(0); tc_fw_src.push_back(0); tc_fw_k.push_back(0); }
    if (tc_bw_col.empty()) { tc_bw_col.push_back(0); tc_bw_tail.push_back(0); tc_bw_src.push_back(0); tc_bw_i.push_back(0); }
    h->tc_nroots = (int)tc_roots.size();
    CK(upload(h->d_tc_bw_ptr, tc_bw_ptr, s)); CK(upload(h->d_tc_bw_col, tc_bw_col, s)); CK(upload(h->d_tc_bw_tail, tc_bw_tail, s)); CK(upload(h->d_tc_bw_src, tc_bw_src, s));
    CK(upload(h->d_tc_bw_i, tc_bw_i, s)); CK(upload(h->d_tc_roots, tc_roots, s));
    CK(upload(h->d_tc_pan_k, tc_pan_k, s)); CK(upload(h->d_tc_pan_x, tc_pan_x, s)); CK(upload(h->d_tc_upd_ptr, tc_upd_ptr, s)); CK(upload(h->d_tc_upd_fac, tc_upd_fac, s)); CK(upload(h->d_tc_upd_dst, tc_upd_dst, s));
    CK(upload(h->d_tc_upd_a, tc_upd_a, s)); CK(upload(h->d_tc_upd_b, tc_upd_b, s)); CK(upload(h->d_tc_fw_ptr, tc_fw_ptr, s)); CK(upload(h->d_tc_fw_row, tc_fw_row, s));
    CK(upload(h->d_tc_fw_tail, tc_fw_tail, s)); CK(upload(h->d_tc_fw_src, tc_fw_src, s)); CK(upload(h->d_tc_fw_k, tc_fw_k, s)); CK(upload(h->d_tc_lev_cols, tc_lev_cols, s));
    CK(upload(h->d_tc_tile_off, tc_tile_off, s)); CK(upload(h->d_tc_col_ptr, h->tc_col_ptr, s)); CK(upload(h->d_tc_col_rows, tc_col_rows, s));
    h->tc_T_bytes = sizeof(double) * kTileDoubles * (size_t)tc_tiles;
    CK(h->d_tc_T.alloc(h->tc_T_bytes)); CK(h->d_tc_W.alloc(sizeof(double) * kTileDoubles * (size_t)nT)); CK(h->d_tc_b.alloc(sizeof(double) * kTS * (size_t)nT));
    h->TC.nT = nT; h->TC.n = nf_dim;
  }
  if (h->pcg) {
    int dev_sms = 0, per_sm = 0;
    CK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, h->device));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pcg_kernel, 256, 0));
    if (per_sm < 1) return h->fail(BSNLS_ERR_CUDA, "pcg_kernel cannot be made resident");
    h->pcg_grid = dev_sms * std::min(per_sm, 2);
    CK(h->d_pcg_x.alloc(sizeof(double) * nf_dim)); CK(h->d_pcg_r.alloc(sizeof(double) * nf_dim)); CK(h->d_pcg_z.alloc(sizeof(double) * nf_dim));
    CK(h->d_pcg_p.alloc(sizeof(double) * nf_dim)); CK(h->d_pcg_Ap.alloc(sizeof(double) * nf_dim)); CK(h->d_pcg_Minv.alloc(sizeof(double) * 9 * (size_t)h->nf));
    CK(h->d_pcg_part.alloc(sizeof(double) * 5 * (size_t)std::max(h->pcg_grid, BlocksFor(h->nf, 8) + 1)));
    if (h->world > 1) {   // block rows of S partitioned over the ranks, balanced by non-zero blocks
      bsnls_shard_range(h->bsr_rowptr.data(), h->nf, h->rank, h->world, &h->pcgd_rb0, &h->pcgd_rb1);   // the rule of the landmark shards (tests/test_sharding_cpu.py)
      CK(h->d_pcgd_zsend.alloc(sizeof(double) * (nf_dim + 3))); CK(h->d_pcgd_zfull.alloc(sizeof(double) * (nf_dim + 3))); CK(h->d_pcgd_xsend.alloc(sizeof(double) * nf_dim));
      CK(h->d_pcgd_scal.alloc(sizeof(double) * 4)); CK(h->d_pcgd_state.alloc(sizeof(PcgDistState)));
      if (!h->pcgd_done_host) CK(cudaMallocHost(&h->pcgd_done_host, sizeof(int)));
    }
  } CK(h->d_poseH.alloc(sizeof(double) * 21 * (size_t)std::max(1, h->np)));
  CK(h->d_gfJ.alloc(sizeof(double) * joff)); CK(h->d_gfR.alloc(sizeof(double) * roff));
  CK(h->d_cost_part.alloc(sizeof(double) * (h->nb_eval + h->ngf + 1))); CK(h->d_cand_part.alloc(sizeof(double) * (h->nb_obs + h->ngf + 1)));
  CK(h->d_model_part.alloc(sizeof(double) * (h->nb_lm + h->ngf + 1)));
  CK(h->d_step_part.alloc(sizeof(double) * h->nb_act)); CK(h->d_candsq_part.alloc(sizeof(double) * h->nb_act));
  CK(h->d_gsq_part.alloc(sizeof(double) * h->nb_act)); CK(h->d_gmax_part.alloc(sizeof(double) * h->nb_act)); CK(h->d_xsq_part.alloc(sizeof(double) * h->nb_act));
  CK(h->d_red.alloc(sizeof(double) * 32)); CK(h->d_state.alloc(sizeof(LMState)));
  phase("work buffers + tile tables");
  CK(h->arena.Commit(s)); g_rec = nullptr;   // pointers are valid from here on
  CK(cudaEventRecord(h->ev_stage, s)); h->stage_busy = true;
  // views into the zeroed block (non-pose entries of gradL / colsqL stay 0)
  h->d_gradL.p = h->d_zero.p; h->d_gradL.view = true; h->d_gradL.bytes = sizeof(double) * nf_dim;
  h->d_colsqL.p = h->d_zero.as<double>() + nf_dim; h->d_colsqL.view = true; h->d_colsqL.bytes = sizeof(double) * nf_dim;
  h->d_y.p = h->d_zero.as<double>() + 2 * (size_t)nf_dim; h->d_y.view = true; h->d_y.bytes = sizeof(double) * ntan;
  CK(cudaMemsetAsync(h->d_zero.p, 0, sizeof(double) * (2 * (size_t)nf_dim + ntan), s));
  if (h->tiled) {
    h->TC.bw_ptr = h->d_tc_bw_ptr.as<int>();
    h->TC.bw_col = h->d_tc_bw_col.as<int>();
    h->TC.bw_tail = h->d_tc_bw_tail.as<int>();
    h->TC.bw_src = h->d_tc_bw_src.as<long long>();
    h->TC.bw_i = h->d_tc_bw_i.as<int>();
    h->TC.roots = h->d_tc_roots.as<int>();
    h->TC.pan_k = h->d_tc_pan_k.as<int>();
    h->TC.pan_x = h->d_tc_pan_x.as<int>();
    h->TC.upd_ptr = h->d_tc_upd_ptr.as<int>();
    h->TC.upd_dst = h->d_tc_upd_dst.as<long long>();
    h->TC.upd_fac = h->d_tc_upd_fac.as<int>();
    h->TC.blk_slot = h->d_tc_blk_slot.as<int>();
    h->TC.slot_blk = h->d_tc_slot_blk.as<int>();
    h->TC.row_ptr = h->d_tc_row_ptr.as<int>(); h->TC.row_tile = h->d_tc_row_tile.as<int>();
    h->TC.braw = h->d_tc_braw.as<double>(); h->TC.P = h->d_tc_P.as<double>();
    h->TC.upd_a = h->d_tc_upd_a.as<long long>();
    h->TC.upd_b = h->d_tc_upd_b.as<long long>();
    h->TC.fw_ptr = h->d_tc_fw_ptr.as<int>();
    h->TC.fw_row = h->d_tc_fw_row.as<int>();
    h->TC.fw_tail = h->d_tc_fw_tail.as<int>();
    h->TC.fw_src = h->d_tc_fw_src.as<long long>();
    h->TC.fw_k = h->d_tc_fw_k.as<int>();
    h->TC.lev_cols = h->d_tc_lev_cols.as<int>();
    h->TC.tile_off = h->d_tc_tile_off.as<long long>();
    h->TC.col_ptr = h->d_tc_col_ptr.as<int>();
    h->TC.col_rows = h->d_tc_col_rows.as<int>();
    h->TC.T = h->d_tc_T.as<double>();
    h->TC.W = h->d_tc_W.as<double>();
    h->TC.b = h->d_tc_b.as<double>();
  }

  DeviceProblem& P = h->P;
  P.num_blocks = nb; P.num_active = (int)h->active_blocks.size(); P.blk_off = h->d_blk_off.as<int>(); P.blk_size = h->d_blk_size.as<int>();
  P.blk_tan = h->d_blk_tan.as<int>(); P.blk_manifold = h->d_blk_manifold.as<int>(); P.active_blocks = h->d_active.as<int>();
  P.ne = h->ne; P.nf = h->nf; P.ne_dim = ne_dim; P.nf_dim = nf_dim; P.ntan = ntan;
  P.num_poses = h->np; P.pose_q_off = h->d_pose_q_off.as<int>(); P.pose_t_off = h->d_pose_t_off.as<int>();
  P.pose_q_tan = h->d_pose_q_tan.as<int>(); P.pose_t_tan = h->d_pose_t_tan.as<int>(); P.pose_ptr = h->d_pose_ptr.as<int>(); P.pose_obs = h->d_pose_obs.as<int>();
  P.num_cams = pr->num_cameras; P.num_obs = h->nobs; P.obs_pose = h->d_obs_pose.as<int>(); P.obs_lm = h->d_obs_lm.as<int>(); P.obs_cam = h->d_obs_cam.as<int>(); P.obs_pixel = h->d_obs_pixel.as<double>();
  P.num_eval_items = h->n_eval_items; P.obs_pm = h->d_obs_pm.as<int>(); P.eval_items = h->d_eval_items.as<int4>(); P.eval_item_ptr = h->d_eval_item_ptr.as<int>();
  P.pm_lm_off = h->d_pm_lm_off.as<int>(); P.pm_cam = h->d_pm_cam.as<int>(); P.pm_pixel = h->d_pm_pixel.as<double>();
  P.lm_ptr = h->d_lm_ptr.as<int>(); P.lm_off = h->d_lm_off.as<int>(); P.cams = h->d_cams.as<CameraModel>();
  P.num_gf = h->ngf; P.gf_type = h->d_gf_type.as<int>(); P.gf_nblk = h->d_gf_nblk.as<int>(); P.gf_nres = h->d_gf_nres.as<int>(); P.gf_blocks = h->d_gf_blocks.as<int>();
  P.gf_coff = h->d_gf_coff.as<long long>(); P.gf_consts = h->d_gf_consts.as<double>(); P.gf_loss_type = h->d_gf_loss_type.as<int>(); P.gf_loss_a = h->d_gf_loss_a.as<double>();
  P.gf_joff = h->d_gf_joff.as<int>(); P.gf_roff = h->d_gf_roff.as<int>();
  P.num_pairs = h->npairs; P.pair_a = h->d_pair_a.as<int>(); P.pair_b = h->d_pair_b.as<int>(); P.pair_ptr = h->d_pair_ptr.as<int>();
  P.item_i = h->d_item_i.as<int>(); P.item_j = h->d_item_j.as<int>();
  P.chunk_max_slots = h->chunk_max_slots; P.chunk_pose = h->d_chunk_pose.as<int>(); P.chunk_stage_ptr = h->d_chunk_stage_ptr.as<int>(); P.stage_lm = h->d_stage_lm.as<int>();
  P.num_chunks = h->nchunks; P.obs_slot = h->d_obs_slot.as<unsigned short>(); P.syrk_plan = h->d_syrk_plan.as<int>(); P.chunk_begin = h->d_chunk_begin.as<int>(); P.chunk_end = h->d_chunk_end.as<int>(); P.chunk_nslots = h->d_chunk_nslots.as<int>();
  P.pp_ptr = h->d_pp_ptr.as<int>();
  P.chunk_tile_off = h->d_chunk_tile_off.as<long long>(); P.pp_toff = h->d_pp_toff.as<long long>(); P.pp_info = h->d_pp_info.as<int>();
  P.fg_ptr = h->d_fg_ptr.as<int>(); P.fg_fac = h->d_fg_fac.as<int>(); P.fg_lb = h->d_fg_lb.as<int>();
  P.num_gdest = h->ngdest; P.gd_row = h->d_gd_row.as<int>(); P.gd_col = h->d_gd_col.as<int>(); P.gd_ptr = h->d_gd_ptr.as<int>();
  P.gd_fac = h->d_gd_fac.as<int>(); P.gd_bi = h->d_gd_bi.as<int>(); P.gd_bj = h->d_gd_bj.as<int>();
  P.sparse = h->sparse ? 1 : 0; P.nnzb = h->nnzb; P.bsr_rowptr = h->d_bsr_rowptr.as<int>(); P.bsr_col = h->d_bsr_col.as<int>();
  P.diag_slot = h->d_diag_slot.as<int>(); P.pair_slot = h->d_pair_slot.as<int>(); P.gd_slot = h->d_gd_slot.as<int>();
  phase("arena commit + copy");   // asynchronous: the next set_problem waits on ev_stage before it reuses the staging buffer
  h->has_problem = true; h->graphs_valid = false; h->graph_rp_valid = false; h->solves_on_structure = 0;
  return BSNLS_OK;
}

int bsnls_set_values(bsnls_handle* h, const double* v) {
  if (!h || !v) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_set_values before bsnls_set_problem");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(h->d_x.p, v, sizeof(double) * h->num_values, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->jac_valid = false;
  return BSNLS_OK;
}

int bsnls_get_values(bsnls_handle* h, double* v) {
  if (!h || !v) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_get_values before bsnls_set_problem");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(v, h->d_x.p, sizeof(double) * h->num_values, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return BSNLS_OK;
}

int bsnls_solve(bsnls_handle* h, bsnls_summary* sum, bsnls_iteration* its, int32_t max_its_out) {
  NvtxRange nvtx_range("bsnls_solve");
  if (!h || !sum) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_solve before bsnls_set_problem");
  CK(cudaSetDevice(h->device));
  const auto t0 = std::chrono::steady_clock::now();
  std::memset(sum, 0, sizeof(*sum));
  const Options o = MakeOptions(h->opt);
  const int max_it = h->opt.max_num_iterations;
  h->max_records = max_it + 2;
  CK(h->d_recs.alloc(sizeof(bsnls_iteration) * h->max_records));
  int rc = InitState(h, false);
  if (rc != BSNLS_OK) return rc;
  LMState* st = h->d_state.as<LMState>();
  bsnls_iteration* recs = h->d_recs.as<bsnls_iteration>();
  CK(cudaMemcpyAsync(h->d_x0.p, h->d_x.p, sizeof(double) * h->num_values, cudaMemcpyDeviceToDevice, h->stream));
  CK(cudaMemcpyAsync(h->d_xc.p, h->d_x.p, sizeof(double) * h->num_values, cudaMemcpyDeviceToDevice, h->stream));  // constant blocks
  CK(cudaEventRecord(h->ev0, h->stream));
  if ((rc = EnqueueXNorm(h)) != BSNLS_OK) return rc;
  // A wall-clock budget (optimizeFor, hash_graph.cpp:428-444) is checked on the host before every pass, as Ceres checks it
  // when it finalises an iteration. Ranks of a sharded solve must issue identical collectives, so they agree on the
  // decision first (max over ranks of the "spent" flag).
  const bool timed = h->opt.max_solver_time_in_seconds < 1e8;
  auto budget_spent = [&](int* spent) -> int {
    *spent = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() >= h->opt.max_solver_time_in_seconds ? 1 : 0;
    if (h->world <= 1) return BSNLS_OK;
    double f = *spent;
    CK(h->d_flag.alloc(sizeof(double)));
    CK(cudaMemcpyAsync(h->d_flag.p, &f, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    const int rc2 = AllReduce(h, h->d_flag.as<double>(), 1, kNcclMax);
    if (rc2 != BSNLS_OK) return rc2;
    CK(cudaMemcpyAsync(&f, h->d_flag.p, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *spent = f > 0.0 ? 1 : 0;
    return BSNLS_OK;
  };
  auto enqueue_tail = [&]() -> int { return EnqueueTail(h, o, recs); };
  auto enqueue_pass = [&]() -> int { return EnqueuePass(h, o, recs); };
  // Small reduced systems are launch-latency bound: replay captured graphs instead of enqueuing kernel by kernel.
  // the launch sequence of a pass is fixed per problem (the cooperative PCG launch is not captured). A structure that is
  // solved once (a sliding window changes every smoother cycle) is launched directly; big reduced systems (hundreds of
  // launches per pass) are captured at once.
  // Sharded solves replay graphs as well (the NCCL collectives are captured with the kernels): at 8 GPUs a pass is a few
  // hundred microseconds of device work behind ~40 launches and two collectives, and enqueuing them one by one is slower
  // than executing them. NCCL sets its connections up lazily, so the first collectives of a communicator run uncaptured.
  const bool graphs = h->use_graphs && !h->pcg && (h->solves_on_structure > 0 || h->tiled) && (h->world == 1 || h->nccl_warm);
  h->solves_on_structure++;
  if (graphs && (!h->graphs_valid || h->graph_recs != recs)) {
    if (h->graph_pass) { cudaGraphExecDestroy(h->graph_pass); h->graph_pass = nullptr; }
    if (h->graph_tail) { cudaGraphExecDestroy(h->graph_tail); h->graph_tail = nullptr; }
    cudaGraph_t g = nullptr;
    const int64_t launches_before = h->launches;
    CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    rc = enqueue_pass();
    cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
    if (rc != BSNLS_OK || ce != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "graph capture (pass): %s", cudaGetErrorString(ce));
    CK(cudaGraphInstantiate(&h->graph_pass, g, 0)); cudaGraphDestroy(g);
    h->graph_pass_launches = h->launches - launches_before;
    CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    rc = enqueue_tail();
    ce = cudaStreamEndCapture(h->stream, &g);
    if (rc != BSNLS_OK || ce != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "graph capture (tail): %s", cudaGetErrorString(ce));
    CK(cudaGraphInstantiate(&h->graph_tail, g, 0)); cudaGraphDestroy(g);
    h->graph_tail_launches = h->launches - launches_before - h->graph_pass_launches; h->launches = launches_before;
    h->graph_recs = recs;
    h->graphs_valid = true;
  }
  for (int pass = 0; pass <= max_it; ++pass) {
    if (timed && pass < max_it) {
      int spent = 0;
      if ((rc = budget_spent(&spent)) != BSNLS_OK) return rc;
      if (spent) {
        // iteration zero, or the Jacobian pass an accepted step still needs for its record; then stop (NO_CONVERGENCE)
        if ((rc = enqueue_tail()) != BSNLS_OK) return rc;
        LAUNCH(h, set_done_kernel, 1, 1, st, (int)BSNLS_NO_CONVERGENCE, (int)MSG_MAX_TIME);
        break;
      }
    }
    if (pass == max_it) {
      if (graphs) { CK(cudaGraphLaunch(h->graph_tail, h->stream)); h->launches += h->graph_tail_launches; } else if ((rc = enqueue_tail()) != BSNLS_OK) return rc;
      break;
    }
    if (graphs) { CK(cudaGraphLaunch(h->graph_pass, h->stream)); h->launches += h->graph_pass_launches; } else if ((rc = enqueue_pass()) != BSNLS_OK) return rc;
    if (timed || (pass & 3) == 3) {
      // poll the device state so that converged solves stop enqueuing (every rank reads the same replicated state)
      LMState hs; CK(cudaMemcpyAsync(&hs, st, sizeof(hs), cudaMemcpyDeviceToHost, h->stream)); CK(cudaStreamSynchronize(h->stream));
      if (hs.done) break;
    }
  }
  if ((rc = GatherValues(h)) != BSNLS_OK) return rc;
  if (h->world > 1) h->nccl_warm = true;
  CK(cudaEventRecord(h->ev1, h->stream));
  LMState hs;
  CK(cudaMemcpyAsync(&hs, st, sizeof(hs), cudaMemcpyDeviceToHost, h->stream));
  std::vector<bsnls_iteration> hrec(h->max_records);
  CK(cudaMemcpyAsync(hrec.data(), recs, sizeof(bsnls_iteration) * h->max_records, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaGetLastError());
  float ms = 0; CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  if (!hs.done) { hs.termination = BSNLS_NO_CONVERGENCE; hs.message_id = MSG_MAX_ITER; }
  sum->termination_type = hs.termination;
  std::snprintf(sum->message, sizeof(sum->message), "%s", MessageText(hs.message_id));
  const int nrec = std::min(hs.num_records, h->max_records);
  sum->num_iterations = nrec; sum->num_successful_steps = hs.num_successful; sum->num_unsuccessful_steps = hs.num_unsuccessful;
  sum->fixed_cost = hs.fixed_cost;
  sum->initial_cost = nrec ? hrec[0].cost : 0.0; sum->final_cost = sum->initial_cost;
  for (int i = 0; i < nrec; ++i) sum->final_cost = std::min(sum->final_cost, hrec[i].cost);  // ceres solver.cc SetSummaryFinalCost
  for (int i = 0; i < nrec && i < max_its_out && its; ++i) its[i] = hrec[i];
  sum->num_e_blocks = h->ne_global; sum->num_f_blocks = h->nf; sum->reduced_system_dim = h->P.nf_dim; sum->linear_solver_used = h->pcg ? BSNLS_LINEAR_ITERATIVE_SCHUR : BSNLS_LINEAR_DENSE_SCHUR;
  { int64_t lower = 0; for (int r = 0; r < h->nf; ++r) for (int k = h->bsr_rowptr[r]; k < h->bsr_rowptr[r + 1]; ++k) lower += h->bsr_col[k] <= r; sum->reduced_system_nnz_blocks = lower; }
  sum->num_parameter_blocks_reduced = h->ne_global + h->nf; sum->num_effective_parameters_reduced = 3 * (h->ne_global + h->nf);
  sum->num_parameters_reduced = 3 * h->ne_global; for (int b : h->fblocks) sum->num_parameters_reduced += h->blk_size[b];
  sum->num_residual_blocks_reduced = h->nobs_global + h->ngf_active;
  sum->num_residuals_reduced = 2 * h->nobs_global; for (int f = 0; f < h->ngf_active; ++f) sum->num_residuals_reduced += h->gf_nres_h[f];
  sum->device_time_in_seconds = 1e-3 * ms;
  if (hs.termination == BSNLS_FAILURE) {  // leave the parameters as they were
    CK(cudaMemcpyAsync(h->d_x.p, h->d_x0.p, sizeof(double) * h->num_values, cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  h->jac_valid = false;
  sum->total_time_in_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return BSNLS_OK;
}

int bsnls_run_eval(bsnls_handle* h, double* cost) {
  if (!h) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_run_eval before bsnls_set_problem");
  CK(cudaSetDevice(h->device));
  int rc = InitState(h, false);
  if (rc != BSNLS_OK) return rc;
  h->profiling = true;
  CK(cudaEventRecord(h->ev0, h->stream));
  EnqueueEval(h);
  CK(cudaEventRecord(h->ev1, h->stream));
  h->profiling = false;
  RedList L{}; L.gate = 0; L.count = 0;
  if (h->nobs) L.d[L.count++] = Desc(h->d_cost_part.as<double>(), h->nb_eval, 0, RED_COST, false);
  if (h->ngf) L.d[L.count++] = Desc(h->d_cost_part.as<double>() + h->nb_eval, h->ngf, 0, RED_COST, true);
  LAUNCH(h, reduce_kernel, 1, 256, L, h->d_red.as<double>(), h->d_state.as<LMState>());
  if ((rc = AllReduceScalars(h)) != BSNLS_OK) return rc;
  double red[kRedSize];
  CK(cudaMemcpyAsync(red, h->d_red.p, sizeof(red), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  CK(cudaEventElapsedTime(&h->last_eval_ms, h->ev0, h->ev1));
  if (cost) *cost = red[RED_COST] + red[RED_NUM + RED_COST];
  cudaEventElapsedTime(&h->breakdown[0], h->pev[0], h->pev[1]); cudaEventElapsedTime(&h->breakdown[1], h->pev[1], h->pev[2]);
  h->jac_valid = true;
  return BSNLS_OK;
}

int bsnls_run_schur(bsnls_handle* h, double radius) {
  if (!h) return BSNLS_ERR_INVALID;
  if (!h->has_problem || !h->jac_valid) return h->fail(BSNLS_ERR_STATE, "bsnls_run_schur needs bsnls_run_eval first");
  if (!(radius > 0)) return h->fail(BSNLS_ERR_INVALID, "radius must be positive");
  CK(cudaSetDevice(h->device));
  LMState s{}; s.radius = radius; s.decrease_factor = 2.0; s.need_jacobian = 1; s.scale_ready = 0;
  CK(cudaMemcpyAsync(h->d_state.p, &s, sizeof(s), cudaMemcpyHostToDevice, h->stream));
  const Options o = MakeOptions(h->opt);
  h->profiling = true;
  CK(cudaEventRecord(h->ev0, h->stream));
  MARK(h, 3);
  EnqueueLandmark(h, o);
  MARK(h, 4);
  int rc = EnqueueLinearise(h, o, true);
  if (rc != BSNLS_OK) { h->profiling = false; return rc; }
  EnqueueAssembleFinish(h, o);
  CK(cudaEventRecord(h->ev1, h->stream));
  h->profiling = false;
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  CK(cudaEventElapsedTime(&h->last_schur_ms, h->ev0, h->ev1));
  cudaEventElapsedTime(&h->breakdown[2], h->pev[3], h->pev[4]); cudaEventElapsedTime(&h->breakdown[3], h->pev[4], h->pev[5]);
  cudaEventElapsedTime(&h->breakdown[4], h->pev[5], h->pev[6]); cudaEventElapsedTime(&h->breakdown[5], h->pev[6], h->pev[7]);
  cudaEventElapsedTime(&h->breakdown[7], h->pev[7], h->pev[9]); cudaEventElapsedTime(&h->breakdown[6], h->pev[9], h->pev[8]);
  return BSNLS_OK;
}

int bsnls_enable_breakdown(bsnls_handle* h, int32_t on) {
  if (!h) return BSNLS_ERR_INVALID;
  h->breakdown_enabled = on != 0;
  return BSNLS_OK;
}

int bsnls_run_pass(bsnls_handle* h, double radius, double* cost) {
  NvtxRange nvtx_range("bsnls_run_pass");
  if (!h) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_run_pass before bsnls_set_problem");
  if (!(radius > 0)) return h->fail(BSNLS_ERR_INVALID, "radius must be positive");
  CK(cudaSetDevice(h->device));
  int rc = InitState(h, false, radius);
  if (rc != BSNLS_OK) return rc;
  const Options o = MakeOptions(h->opt);
  auto enqueue = [&]() -> int {
    EnqueueEval(h, 1);
    MARK(h, 3);
    if (!h->fused) EnqueueLandmark(h, o);
    MARK(h, 4);
    const int r2 = EnqueueLinearise(h, o, true);
    if (r2 != BSNLS_OK) return r2;
    EnqueueAssembleFinish(h, o);
    return BSNLS_OK;
  };
  // One graph launch per pass unless the per-kernel breakdown is wanted (events cannot be timed inside a graph). The first
  // pass of a sharded handle runs directly: NCCL connects lazily.
  const bool graph = h->use_graphs && !h->breakdown_enabled && (h->world == 1 || h->nccl_warm);
  if (graph && !h->graph_rp_valid) {
    if (h->graph_rp) { cudaGraphExecDestroy(h->graph_rp); h->graph_rp = nullptr; }
    cudaGraph_t g = nullptr;
    const int64_t before = h->launches;
    CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    rc = enqueue();
    const cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
    if (rc != BSNLS_OK || ce != cudaSuccess) return h->fail(BSNLS_ERR_CUDA, "graph capture (run_pass): %s", cudaGetErrorString(ce));
    CK(cudaGraphInstantiate(&h->graph_rp, g, 0)); cudaGraphDestroy(g);
    h->graph_rp_launches = h->launches - before; h->launches = before;
    h->graph_rp_valid = true;
  }
  h->profiling = !graph;
  CK(cudaEventRecord(h->ev0, h->stream));
  if (graph) { CK(cudaGraphLaunch(h->graph_rp, h->stream)); h->launches += h->graph_rp_launches; }
  else if ((rc = enqueue()) != BSNLS_OK) { h->profiling = false; return rc; }
  CK(cudaEventRecord(h->ev1, h->stream));
  h->profiling = false;
  if (h->world > 1) h->nccl_warm = true;
  // cost = what the gradient pass reduced (local + replicated parts)
  double red[kRedSize];
  CK(cudaMemcpyAsync(red, h->d_red.p, sizeof(red), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  if (cost) *cost = red[RED_COST] + red[RED_NUM + RED_COST];
  float ms = 0; CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  h->last_eval_ms = ms; h->last_schur_ms = 0.f;
  if (!graph) {
    cudaEventElapsedTime(&h->breakdown[0], h->pev[0], h->pev[1]); cudaEventElapsedTime(&h->breakdown[1], h->pev[1], h->pev[2]);
    cudaEventElapsedTime(&h->breakdown[2], h->pev[3], h->pev[4]); cudaEventElapsedTime(&h->breakdown[3], h->pev[4], h->pev[5]);
    cudaEventElapsedTime(&h->breakdown[4], h->pev[5], h->pev[6]); cudaEventElapsedTime(&h->breakdown[5], h->pev[6], h->pev[7]);
    cudaEventElapsedTime(&h->breakdown[7], h->pev[7], h->pev[9]); cudaEventElapsedTime(&h->breakdown[6], h->pev[9], h->pev[8]);
  }
  h->jac_valid = true;
  return BSNLS_OK;
}

int bsnls_last_kernel_ms(bsnls_handle* h, float* eval_ms, float* schur_ms) {
  if (!h) return BSNLS_ERR_INVALID;
  if (eval_ms) *eval_ms = h->last_eval_ms;
  if (schur_ms) *schur_ms = h->last_schur_ms;
  return BSNLS_OK;
}

int bsnls_last_breakdown_ms(bsnls_handle* h, float out[8]) {
  if (!h || !out) return BSNLS_ERR_INVALID;
  std::memcpy(out, h->breakdown, sizeof(float) * 8);
  return BSNLS_OK;
}

int bsnls_get_reduced_system(bsnls_handle* h, double radius, double* S, double* rhs, int32_t dim) {
  if (!h || !S || !rhs) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "no problem");
  if (dim != h->P.nf_dim) return h->fail(BSNLS_ERR_INVALID, "dim must be %d", h->P.nf_dim);
  int rc = bsnls_run_eval(h, nullptr);
  if (rc != BSNLS_OK) return rc;
  rc = bsnls_run_schur(h, radius);
  if (rc != BSNLS_OK) return rc;
  CK(cudaMemcpyAsync(rhs, h->d_rhs.p, sizeof(double) * dim, cudaMemcpyDeviceToHost, h->stream));
  if (h->sparse) {
    std::vector<double> blocks((size_t)h->nnzb * 9);
    CK(cudaMemcpyAsync(blocks.data(), h->d_S.p, sizeof(double) * blocks.size(), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    std::memset(S, 0, sizeof(double) * (size_t)dim * dim);
    // internal (bandwidth-reducing) f-block order -> API (program) order
    for (int r = 0; r < h->nf; ++r)
      for (int k = h->bsr_rowptr[r]; k < h->bsr_rowptr[r + 1]; ++k)
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) S[(size_t)(3 * h->f_api[h->bsr_col[k]] + j) * dim + 3 * h->f_api[r] + i] = blocks[(size_t)k * 9 + 3 * i + j];
    std::vector<double> tmp(rhs, rhs + dim);
    for (int r = 0; r < h->nf; ++r) for (int i = 0; i < 3; ++i) rhs[3 * h->f_api[r] + i] = tmp[3 * r + i];
    return BSNLS_OK;
  }
  CK(cudaMemcpyAsync(S, h->d_S.p, sizeof(double) * (size_t)dim * dim, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  for (int c = 0; c < dim; ++c) for (int r = 0; r < c; ++r) S[(size_t)c * dim + r] = S[(size_t)r * dim + c];  // mirror the lower triangle
  return BSNLS_OK;
}

int bsnls_get_reduced_blocks(bsnls_handle* h, double radius, int32_t* rows, int32_t* cols, double* vals, int64_t* count, double* rhs) {
  if (!h || !count) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "no problem");
  int64_t n = 0;
  int rc = bsnls_get_reduced_pattern(h, nullptr, nullptr, &n);
  if (rc != BSNLS_OK) return rc;
  if (!rows || !cols || !vals) { *count = n; return BSNLS_OK; }
  if (*count < n) return h->fail(BSNLS_ERR_INVALID, "block buffers too small");
  if ((rc = bsnls_get_reduced_pattern(h, rows, cols, count)) != BSNLS_OK) return rc;
  if ((rc = bsnls_run_eval(h, nullptr)) != BSNLS_OK) return rc;
  if ((rc = bsnls_run_schur(h, radius)) != BSNLS_OK) return rc;
  const int dim = h->P.nf_dim;
  std::vector<double> r(dim);
  CK(cudaMemcpyAsync(r.data(), h->d_rhs.p, sizeof(double) * dim, cudaMemcpyDeviceToHost, h->stream));
  std::memset(vals, 0, sizeof(double) * 9 * (size_t)n);
  auto find = [&](int ar, int ac) -> int64_t {   // (row, col) pairs are sorted
    int64_t lo = 0, hi = n;
    while (lo < hi) { const int64_t m = (lo + hi) / 2; if (rows[m] < ar || (rows[m] == ar && cols[m] < ac)) lo = m + 1; else hi = m; }
    return (lo < n && rows[lo] == ar && cols[lo] == ac) ? lo : -1;
  };
  if (h->sparse) {
    std::vector<double> blocks((size_t)h->nnzb * 9);
    CK(cudaMemcpyAsync(blocks.data(), h->d_S.p, sizeof(double) * blocks.size(), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int br = 0; br < h->nf; ++br)
      for (int k = h->bsr_rowptr[br]; k < h->bsr_rowptr[br + 1]; ++k) {
        const int ar = h->f_api[br], ac = h->f_api[h->bsr_col[k]];
        if (ar < ac) continue;
        const int64_t q = find(ar, ac);
        if (q < 0) return h->fail(BSNLS_ERR_STATE, "reduced-system block outside the reported pattern");
        std::memcpy(vals + 9 * q, blocks.data() + (size_t)k * 9, sizeof(double) * 9);
      }
    if (rhs) for (int br = 0; br < h->nf; ++br) for (int i = 0; i < 3; ++i) rhs[3 * h->f_api[br] + i] = r[3 * br + i];
    return BSNLS_OK;
  }
  std::vector<double> S((size_t)dim * dim);
  CK(cudaMemcpyAsync(S.data(), h->d_S.p, sizeof(double) * S.size(), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  for (int64_t q = 0; q < n; ++q)
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
      const int rr = 3 * rows[q] + i, cc = 3 * cols[q] + j;
      vals[9 * q + 3 * i + j] = rr >= cc ? S[(size_t)cc * dim + rr] : S[(size_t)rr * dim + cc];   // lower triangle, column-major
    }
  if (rhs) std::memcpy(rhs, r.data(), sizeof(double) * dim);
  return BSNLS_OK;
}

// replaces HashGraph::getCovariance (hash_graph.cpp:303-414) for pairs of non-landmark blocks: columns of the inverse of the
// undamped reduced system (= the f-f part of (J^T J)^-1), one triangular solve per column on the factor of one
// linearisation.
int bsnls_get_covariance(bsnls_handle* h, const int32_t* pairs, int32_t num_pairs, int32_t use_tangent_space, double* out) {
  NvtxRange nvtx_range("bsnls_get_covariance");
  if (!h || (num_pairs > 0 && (!pairs || !out)) || num_pairs < 0) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_get_covariance before bsnls_set_problem");
  if (num_pairs == 0) return BSNLS_OK;
  if (h->pcg) return h->fail(BSNLS_ERR_UNSUPPORTED, "covariance needs the direct solver (the iterative solver keeps no factor)");
  CK(cudaSetDevice(h->device));
  const int ne_dim = h->P.ne_dim, nf_dim = h->P.nf_dim;
  std::vector<int> col_of(h->num_blocks, -1);     // requested block -> slot in the list of solved block columns
  std::vector<int> cols;
  for (int q = 0; q < 2 * num_pairs; ++q) {
    const int b = pairs[q];
    if (b < 0 || b >= h->num_blocks) return h->fail(BSNLS_ERR_INVALID, "covariance request: block index out of range");
    if (h->blk_tan[b] < 0) return h->fail(BSNLS_ERR_UNSUPPORTED, "covariance request: block %d is constant or unused (ceres::Covariance has no block for it)", b);
    if (h->blk_tan[b] < ne_dim) return h->fail(BSNLS_ERR_UNSUPPORTED, "covariance request: block %d is an eliminated landmark; only blocks of the reduced system are supported", b);
    if (col_of[b] < 0 && (q & 1)) { col_of[b] = (int)cols.size(); cols.push_back(b); }
  }
  // linearise at the current values without damping: radius -> infinity makes D^2 vanish
  int rc = InitState(h, false, 1e300);
  if (rc != BSNLS_OK) return rc;
  const Options o = MakeOptions(h->opt);
  LMState* st = h->d_state.as<LMState>();
  EnqueueEval(h, 1);
  if (!h->fused) EnqueueLandmark(h, o);
  if ((rc = EnqueueLinearise(h, o, true)) != BSNLS_OK) return rc;
  EnqueueAssembleFinish(h, o);
  std::vector<double> sol((size_t)cols.size() * 3 * nf_dim);
  for (size_t c = 0; c < cols.size(); ++c)
    for (int k = 0; k < 3; ++k) {
      const int j = h->blk_tan[cols[c]] - ne_dim + k;
      LAUNCH(h, unit_rhs_kernel, BlocksFor(nf_dim, 256), 256, h->d_rhs.as<double>(), nf_dim, j);
      EnqueueDenseSolve(h, c == 0 && k == 0);
      CK(cudaMemcpyAsync(sol.data() + (c * 3 + k) * nf_dim, h->d_yf.p, sizeof(double) * nf_dim, cudaMemcpyDeviceToHost, h->stream));
    }
  LMState hs;
  CK(cudaMemcpyAsync(&hs, st, sizeof(hs), cudaMemcpyDeviceToHost, h->stream));
  std::vector<double> x;
  if (!use_tangent_space) { x.resize(h->num_values); CK(cudaMemcpyAsync(x.data(), h->d_x.p, sizeof(double) * h->num_values, cudaMemcpyDeviceToHost, h->stream)); }
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  h->jac_valid = false;
  if (hs.lin_fail) return h->fail(BSNLS_ERR_STATE, "the normal equations are rank deficient at the current values: no covariance (ceres::Covariance::Compute returns false)");
  auto plus_jacobian = [&](int b, double* J /* size x 3 */) {   // ambient = J tangent
    if (h->blk_manifold[b] == BSNLS_MANIFOLD_QUATERNION) {
      const double* q = x.data() + h->blk_off[b];
      const double v[12] = {-q[1], -q[2], -q[3], q[0], -q[3], q[2], q[3], q[0], -q[1], -q[2], q[1], q[0]};
      for (int e = 0; e < 12; ++e) J[e] = 0.5 * v[e];
    } else { for (int e = 0; e < 9; ++e) J[e] = (e % 4 == 0) ? 1.0 : 0.0; }
  };
  for (int q = 0; q < num_pairs; ++q) {
    const int bi = pairs[2 * q], bj = pairs[2 * q + 1];
    const int ri = h->blk_tan[bi] - ne_dim, cj = col_of[bj];
    double T[9];
    for (int a = 0; a < 3; ++a) for (int k = 0; k < 3; ++k) T[3 * a + k] = sol[((size_t)cj * 3 + k) * nf_dim + ri + a];
    double* dst = out + 16 * (size_t)q;
    std::memset(dst, 0, sizeof(double) * 16);
    if (use_tangent_space) { std::memcpy(dst, T, sizeof(T)); continue; }
    const int si = h->blk_size[bi], sj = h->blk_size[bj];
    double Ji[12], Jj[12], M[12];
    plus_jacobian(bi, Ji); plus_jacobian(bj, Jj);
    for (int a = 0; a < si; ++a) for (int k = 0; k < 3; ++k) { double acc = 0; for (int m = 0; m < 3; ++m) acc += Ji[3 * a + m] * T[3 * m + k]; M[3 * a + k] = acc; }
    for (int a = 0; a < si; ++a) for (int c = 0; c < sj; ++c) { double acc = 0; for (int k = 0; k < 3; ++k) acc += M[3 * a + k] * Jj[3 * c + k]; dst[a * sj + c] = acc; }
  }
  return BSNLS_OK;
}

int bsnls_trial_step(bsnls_handle* h, const double* values, double radius, int32_t keep_jacobi_scale, bsnls_iteration* out, int32_t* num_out) {
  NvtxRange nvtx_range("bsnls_trial_step");
  if (!h || !out || !num_out) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_trial_step before bsnls_set_problem");
  if (!(radius > 0)) return h->fail(BSNLS_ERR_INVALID, "radius must be positive");
  CK(cudaSetDevice(h->device));
  if (values) CK(cudaMemcpyAsync(h->d_x.p, values, sizeof(double) * h->num_values, cudaMemcpyHostToDevice, h->stream));
  const Options o = MakeOptions(h->opt);
  h->max_records = std::max(h->max_records, 4);
  CK(h->d_recs.alloc(sizeof(bsnls_iteration) * h->max_records));
  int rc = InitState(h, keep_jacobi_scale != 0, radius);
  if (rc != BSNLS_OK) return rc;
  bsnls_iteration* recs = h->d_recs.as<bsnls_iteration>();
  CK(cudaMemcpyAsync(h->d_xc.p, h->d_x.p, sizeof(double) * h->num_values, cudaMemcpyDeviceToDevice, h->stream));
  if ((rc = EnqueueXNorm(h)) != BSNLS_OK) return rc;
  if ((rc = EnqueuePass(h, o, recs)) != BSNLS_OK) return rc;
  if ((rc = EnqueueTail(h, o, recs)) != BSNLS_OK) return rc;
  LMState hs; bsnls_iteration hrec[2];
  CK(cudaMemcpyAsync(&hs, h->d_state.p, sizeof(hs), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(hrec, recs, sizeof(hrec), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  *num_out = std::min(hs.num_records, 2);
  for (int i = 0; i < *num_out; ++i) out[i] = hrec[i];
  h->jac_valid = false;
  return BSNLS_OK;
}

int bsnls_evaluate(bsnls_handle* h, double* cost, double* residuals, int64_t* num_residuals, double* gradient, int64_t* num_gradient) {
  NvtxRange nvtx_range("bsnls_evaluate");
  if (!h) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_evaluate before bsnls_set_problem");
  CK(cudaSetDevice(h->device));
  if (num_residuals) *num_residuals = h->num_res_rows;
  if (num_gradient) *num_gradient = h->P.ntan;
  int rc = InitState(h, false);
  if (rc != BSNLS_OK) return rc;
  const Options o = MakeOptions(h->opt);
  LMState* st = h->d_state.as<LMState>();
  if (h->world > 1) return h->fail(BSNLS_ERR_UNSUPPORTED, "bsnls_evaluate is a single-GPU call (residuals live on the rank that owns their landmark)");
  EnqueueEval(h); if (!h->fused) EnqueueLandmark(h, o); rc = EnqueueLinearise(h, o, false);
  if (rc != BSNLS_OK) return rc;
  // include factors whose blocks are all constant in the cost (Problem::Evaluate evaluates every residual block)
  RedList L{}; L.gate = 0; L.count = 0;
  if (h->nobs) L.d[L.count++] = Desc(h->d_cost_part.as<double>(), h->nb_eval, 0, RED_COST, false);
  if (h->ngf) L.d[L.count++] = Desc(h->d_cost_part.as<double>() + h->nb_eval, h->ngf, 0, RED_COST, true);
  LAUNCH(h, reduce_kernel, 1, 256, L, h->d_red.as<double>(), st);
  double red[kRedSize];
  CK(cudaMemcpyAsync(red, h->d_red.p, sizeof(red), cudaMemcpyDeviceToHost, h->stream));
  std::vector<double> jrec, gfr, grad;
  if (residuals) {
    jrec.resize((size_t)h->nobs * kJRec); gfr.resize(h->gfR_size);
    if (h->nobs) CK(cudaMemcpyAsync(jrec.data(), h->d_Jrec.p, sizeof(double) * jrec.size(), cudaMemcpyDeviceToHost, h->stream));
    if (h->gfR_size) CK(cudaMemcpyAsync(gfr.data(), h->d_gfR.p, sizeof(double) * gfr.size(), cudaMemcpyDeviceToHost, h->stream));
  }
  if (gradient) { grad.resize(h->P.ntan); CK(cudaMemcpyAsync(grad.data(), h->d_grad.p, sizeof(double) * grad.size(), cudaMemcpyDeviceToHost, h->stream)); }
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  if (cost) *cost = red[RED_COST] + red[RED_NUM + RED_COST];
  if (residuals) {
    const std::vector<int>& gf_roff = h->gf_roff_h;
    for (const auto& r : h->res_order) {
      if (r.type == BSNLS_FACTOR_REPROJECTION) { residuals[r.row] = jrec[(size_t)h->obs_pm_h[r.idx] * kJRec + 6]; residuals[r.row + 1] = jrec[(size_t)h->obs_pm_h[r.idx] * kJRec + 7]; }
      else for (int k = 0; k < kNumRes[r.type]; ++k) residuals[r.row + k] = gfr[gf_roff[r.idx] + k];
    }
  }
  if (gradient) {  // tangent gradient over active blocks in program order
    int64_t n = 0;
    for (int b = 0; b < h->num_blocks; ++b) { if (h->blk_tan[b] < 0) continue; for (int k = 0; k < 3; ++k) gradient[n + k] = grad[h->blk_tan[b] + k]; n += 3; }
  }
  h->jac_valid = true;
  return BSNLS_OK;
}

int bsnls_linearize(bsnls_handle* h, int32_t count, const int32_t* residual_blocks, double* residuals, double* jacobians) {
  NvtxRange nvtx_range("bsnls_linearize");
  if (!h) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "bsnls_linearize before bsnls_set_problem");
  if (count < 0 || (count > 0 && (!residual_blocks || !residuals || !jacobians))) return h->fail(BSNLS_ERR_INVALID, "bsnls_linearize: null argument");
  if (h->world > 1) return h->fail(BSNLS_ERR_UNSUPPORTED, "bsnls_linearize is a single-GPU call");
  for (int q = 0; q < count; ++q) if (residual_blocks[q] < 0 || residual_blocks[q] >= (int)h->res_order.size()) return h->fail(BSNLS_ERR_INVALID, "bsnls_linearize: residual block %d out of range", residual_blocks[q]);
  CK(cudaSetDevice(h->device));
  int rc = InitState(h, false);
  if (rc != BSNLS_OK) return rc;
  EnqueueEval(h);     // residuals and Jacobians of every factor at the current values, loss correction applied
  std::vector<double> jrec((size_t)h->nobs * kJRec), gfr(h->gfR_size), gfj(h->gfJ_size);
  if (h->nobs) CK(cudaMemcpyAsync(jrec.data(), h->d_Jrec.p, sizeof(double) * jrec.size(), cudaMemcpyDeviceToHost, h->stream));
  if (h->gfR_size) CK(cudaMemcpyAsync(gfr.data(), h->d_gfR.p, sizeof(double) * gfr.size(), cudaMemcpyDeviceToHost, h->stream));
  if (h->gfJ_size) CK(cudaMemcpyAsync(gfj.data(), h->d_gfJ.p, sizeof(double) * gfj.size(), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError());
  constexpr int kR = BSNLS_MARGINAL_MAX_ROWS, kC = 30;
  std::memset(residuals, 0, sizeof(double) * kR * (size_t)count); std::memset(jacobians, 0, sizeof(double) * kR * kC * (size_t)count);
  for (int q = 0; q < count; ++q) {
    const auto& r = h->res_order[residual_blocks[q]];
    double* R = residuals + (size_t)q * kR; double* J = jacobians + (size_t)q * kR * kC;
    if (r.type == BSNLS_FACTOR_REPROJECTION) {   // record: E[6] | r[2] | Fq[6] | ts; blocks q, t, landmark; F_t = ts E
      const double* rec = jrec.data() + (size_t)h->obs_pm_h[r.idx] * kJRec;
      for (int i = 0; i < 2; ++i) {
        R[i] = rec[6 + i];
        for (int k = 0; k < 3; ++k) { J[i * kC + k] = rec[8 + 3 * i + k]; J[i * kC + 3 + k] = rec[14] * rec[3 * i + k]; J[i * kC + 6 + k] = rec[3 * i + k]; }
      }
    } else {
      const int f = r.idx, nr = h->gf_nres_h[f], ncol = 3 * h->gf_nblk_h[f];
      for (int i = 0; i < nr; ++i) {
        R[i] = gfr[h->gf_roff_h[f] + i];
        for (int k = 0; k < ncol; ++k) J[i * kC + k] = gfj[(size_t)h->gf_joff_h[f] + (size_t)i * ncol + k];
      }
    }
  }
  h->jac_valid = false;
  return BSNLS_OK;
}

int bsnls_get_ordering(bsnls_handle* h, int32_t* e_blocks, int32_t* num_e, int32_t* f_blocks, int32_t* num_f) {
  if (!h) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "no problem");
  if (num_e) *num_e = h->ne_global;
  if (num_f) *num_f = h->nf;
  if (e_blocks) std::memcpy(e_blocks, h->eblocks.data(), sizeof(int) * h->ne_global);
  if (f_blocks) std::memcpy(f_blocks, h->fblocks.data(), sizeof(int) * h->nf);
  return BSNLS_OK;
}

int bsnls_get_reduced_pattern(bsnls_handle* h, int32_t* rows, int32_t* cols, int64_t* count) {
  if (!h || !count) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "no problem");
  std::vector<uint64_t> keys;
  std::vector<int> fpos(h->num_blocks, -1);
  for (int k = 0; k < h->nf; ++k) fpos[h->fblocks[k]] = k;
  auto add = [&](int r, int c) { if (r < 0 || c < 0) return; if (r < c) std::swap(r, c); keys.push_back((uint64_t)r * h->nf + c); };
  for (int p = 0; p < h->npairs; ++p) {
    const int a = h->pair_a[p], b = h->pair_b[p];
    const int fa[2] = {fpos[h->pose_q_blk[a]], fpos[h->pose_t_blk[a]]}, fb[2] = {fpos[h->pose_q_blk[b]], fpos[h->pose_t_blk[b]]};
    for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) add(fa[i], fb[j]);
  }
  for (int d = 0; d < h->ngdest; ++d) add(h->f_api[h->gd_row[d]], h->f_api[h->gd_col[d]]);   // gd_* are internal positions
  for (int k = 0; k < h->nf; ++k) add(k, k);
  std::sort(keys.begin(), keys.end()); keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
  if (rows && cols) { if (*count < (int64_t)keys.size()) return h->fail(BSNLS_ERR_INVALID, "pattern buffers too small"); for (size_t k = 0; k < keys.size(); ++k) { rows[k] = (int)(keys[k] / h->nf); cols[k] = (int)(keys[k] % h->nf); } }
  *count = (int64_t)keys.size();
  return BSNLS_OK;
}

// Landmark shard of one rank: contiguous range [lo, hi) of the elimination order whose observation count is closest
// to an equal share. lm_ptr = prefix sums of observations per landmark (num_landmarks + 1 entries). Pure host logic.
int bsnls_shard_range(const int32_t* lm_ptr, int32_t num_landmarks, int32_t rank, int32_t world_size, int32_t* lo, int32_t* hi) {
  if (!lm_ptr || !lo || !hi || num_landmarks < 0 || world_size < 1 || rank < 0 || rank >= world_size) return BSNLS_ERR_INVALID;
  const long long total = lm_ptr[num_landmarks];
  auto cut = [&](int r) {
    if (r <= 0) return 0;
    if (r >= world_size) return (int)num_landmarks;
    const long long target = total * r / world_size;
    return (int)(std::lower_bound(lm_ptr, lm_ptr + num_landmarks + 1, (int32_t)target) - lm_ptr);
  };
  *lo = std::min<int>(num_landmarks, cut(rank)); *hi = std::max<int>(*lo, std::min<int>(num_landmarks, cut(rank + 1)));
  return BSNLS_OK;
}

int bsnls_get_shard(bsnls_handle* h, int32_t* e_positions, int32_t* count) {
  if (!h || !count) return BSNLS_ERR_INVALID;
  if (!h->has_problem) return h->fail(BSNLS_ERR_STATE, "no problem");
  if (e_positions) { if (*count < (int)h->shard_e.size()) return h->fail(BSNLS_ERR_INVALID, "buffer too small"); std::memcpy(e_positions, h->shard_e.data(), sizeof(int) * h->shard_e.size()); }
  *count = (int32_t)h->shard_e.size();
  return BSNLS_OK;
}

int bsnls_nccl_unique_id(char id[128]) {
  if (!id) return BSNLS_ERR_INVALID;
  if (!Nccl().Load()) { g_create_error = Nccl().error; return BSNLS_ERR_CUDA; }
  NcclUniqueId u;
  const int rc = Nccl().GetUniqueId(&u);
  if (rc != kNcclSuccess) { g_create_error = std::string("ncclGetUniqueId: ") + Nccl().GetErrorString(rc); return BSNLS_ERR_CUDA; }
  std::memcpy(id, u.internal, 128);
  return BSNLS_OK;
}
int bsnls_create_sharded(const bsnls_options* options, int32_t rank, int32_t world_size, const char nccl_id[128], bsnls_handle** handle) {
  if (world_size < 1 || rank < 0 || rank >= world_size) { g_create_error = "bad rank / world_size"; return BSNLS_ERR_INVALID; }
  int rc = bsnls_create(options, handle);
  if (rc != BSNLS_OK || world_size == 1) return rc;
  bsnls_handle* h = *handle;
  if (!nccl_id || !Nccl().Load()) { g_create_error = nccl_id ? Nccl().error : "nccl_id is required"; bsnls_destroy(h); *handle = nullptr; return BSNLS_ERR_CUDA; }
  NcclUniqueId u; std::memcpy(u.internal, nccl_id, 128);
  const int nrc = Nccl().CommInitRank(&h->comm, world_size, u, rank);   // collective over all ranks
  if (nrc != kNcclSuccess) { g_create_error = std::string("ncclCommInitRank: ") + Nccl().GetErrorString(nrc); bsnls_destroy(h); *handle = nullptr; return BSNLS_ERR_CUDA; }
  h->rank = rank; h->world = world_size;
  return BSNLS_OK;
}

}  // extern "C"
