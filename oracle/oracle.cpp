// ORACLE — TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library; the product (beam_slam_b200/) never does.
//
// Restates, on the CPU, what happens behind fuse_graphs::HashGraph::optimize
// (dependencies/fuse/fuse_graphs/src/hash_graph.cpp:416-426): ceres::Problem construction semantics
// (constant blocks removed), residual-block evaluation with manifold + robust-loss correction
// (fuse's own restatement: dependencies/fuse/fuse_constraints/src/marginalize_variables.cpp:281-399),
// Ceres' TrustRegionMinimizer + LevenbergMarquardtStrategy + SchurEliminator with a Cholesky of the reduced system
// (dense for windows, block-sparse with a sparse LL^T for the refinement-sized graphs: oracle_sparse.h)
// (Ceres 1.14/2.0 public sources, third-party, not vendored).
// Elimination ordering = explicit two groups (block_group 0 first), program order inside each group.
//
// Parity status: factor, manifold and linearisation math pinned to the reference's golden vectors
// (tests/test_oracle_golden.py). Iteration-level LM behaviour: PARITY UNPINNED (no Ceres binary here).
#include <omp.h>

#include <chrono>
#include <cstdio>
#include <string>
#include <vector>

#include "oracle_factors.h"
#include "oracle_sparse.h"

namespace orc {

static int NumBlocks(int type) {
  static const int n[BSNLS_NUM_FACTOR_TYPES] = {3, 10, 5, 4, 6, 2, 2, 1, 2, BSNLS_MARGINAL_MAX_BLOCKS, 1, 1};   // block slots per factor
  return n[type];
}
static int NumRes(int type) {
  static const int n[BSNLS_NUM_FACTOR_TYPES] = {2, 15, 15, 6, 6, 6, 3, 3, 3, BSNLS_MARGINAL_MAX_ROWS, 2, 3};
  return n[type];
}
static int ConstStride(int type) {
  static const int n[BSNLS_NUM_FACTOR_TYPES] = {BSNLS_REPROJECTION_STRIDE, BSNLS_IMU_DELTA_STRIDE,
      BSNLS_IMU_PRIOR_STRIDE, BSNLS_POSE_STRIDE, BSNLS_POSE_STRIDE, BSNLS_POSE_STRIDE,
      BSNLS_ORIENTATION_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_VEC3_STRIDE, BSNLS_MARGINAL_STRIDE, BSNLS_GRAVITY_STRIDE, BSNLS_ORIENTATION_STRIDE};
  return n[type];
}

// CostFunction::Evaluate for one factor: residual + ambient row-major Jacobians (jac[b] may be null).
static void EvaluateCost(const bsnls_problem* P, int type, const double* consts, int aux,
                         const double* const* params, double* residual, double** jac) {
  switch (type) {
    case BSNLS_FACTOR_REPROJECTION: {
      const double* cam = P->cameras + (int64_t)aux * BSNLS_CAMERA_STRIDE;
      ReprojectionEvaluate(cam, consts, params[0], params[1], params[2], residual,
                           jac ? jac[0] : nullptr, jac ? jac[1] : nullptr, jac ? jac[2] : nullptr);
      break; }
    case BSNLS_FACTOR_IMU_DELTA: AutoDiffEvaluate<ImuDeltaFunctor, 32>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_IMU_PRIOR: AutoDiffEvaluate<ImuPriorFunctor, 16>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_RELATIVE_POSE: AutoDiffEvaluate<RelPoseFunctor, 14>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_RELATIVE_POSE_EXTRINSICS:
      AutoDiffEvaluate<RelPoseExtFunctor, 21>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_ABSOLUTE_POSE: AutoDiffEvaluate<AbsPoseFunctor, 7>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_RELATIVE_ORIENTATION:
      AutoDiffEvaluate<RelOrientationFunctor, 8>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_ABSOLUTE_VEC3: AutoDiffEvaluate<AbsVec3Functor, 3>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_RELATIVE_VEC3: AutoDiffEvaluate<RelVec3Functor, 6>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_MARGINAL: MarginalEvaluate(consts, params, residual, jac); break;
    case BSNLS_FACTOR_GRAVITY_ALIGNMENT: AutoDiffEvaluate<GravityAlignmentFunctor, 4>(consts, params, residual, jac); break;
    case BSNLS_FACTOR_ABSOLUTE_ORIENTATION: AutoDiffEvaluate<AbsOrientationFunctor, 4>(consts, params, residual, jac); break;
  }
}

struct ResBlock {
  int type, idx;        // group type, index within group
  int nb, nr;           // blocks, residual rows
  int blocks[10];
  int64_t row;          // first row in the residual vector (program order, all residual blocks)
  int64_t jac;          // offset into the tangent Jacobian storage
  int jcols[10];        // tangent column offset inside this block's row-major [nr x ncols] Jacobian, -1 const
  int ncols;
  bool removed;         // all parameter blocks constant -> contributes to fixed_cost only
  int eblock;           // position in e-list of its (single) e-block, or -1
};

struct Model {
  const bsnls_problem* P;
  std::vector<int64_t> off;      // ambient offset per block
  std::vector<int> lsize;        // tangent size per block
  std::vector<char> active;      // non-constant and referenced by a kept residual block
  std::vector<int> tan;          // tangent offset per active block in [e-blocks | f-blocks] order, else -1
  std::vector<int> eblocks, fblocks;
  std::vector<int> epos;         // block -> position in eblocks or -1
  int ne_dim = 0, nf_dim = 0, ntan = 0;
  std::vector<ResBlock> rb;      // program order
  std::vector<std::vector<int>> e_rows;  // per e-block: residual blocks (program order)
  std::vector<int> f_rows;       // residual blocks without e-block
  int64_t nrows = 0, njac = 0;
  std::vector<double> x;         // all blocks, ambient
  double fixed_cost = 0.0;
  std::string error;
};

static const bsnls_factor_group* FindGroup(const bsnls_problem* P, int type) {
  for (int g = 0; g < P->num_groups; ++g) if (P->groups[g].type == type) return &P->groups[g];
  return nullptr;
}

static void EvalResBlockRaw(const Model& M, const ResBlock& r, const double* x, double* res, double** jac) {
  const bsnls_factor_group* g = FindGroup(M.P, r.type);
  const double* params[10];
  for (int b = 0; b < r.nb; ++b) params[b] = x + M.off[r.blocks[b]];
  EvaluateCost(M.P, r.type, g->consts + (int64_t)r.idx * ConstStride(r.type), g->aux ? g->aux[r.idx] : 0,
               params, res, jac);
}
static void LossOf(const Model& M, const ResBlock& r, int* type, double* a) {
  const bsnls_factor_group* g = FindGroup(M.P, r.type);
  if (r.type == BSNLS_FACTOR_REPROJECTION) {
    const double* cam = M.P->cameras + (int64_t)g->aux[r.idx] * BSNLS_CAMERA_STRIDE;
    *type = (int)cam[17]; *a = cam[18];
  } else {
    *type = g->loss_type ? g->loss_type[r.idx] : BSNLS_LOSS_NONE;
    *a = g->loss_param ? g->loss_param[r.idx] : 0.0;
  }
}

static bool BuildModel(const bsnls_problem* P, Model* M) {
  M->P = P;
  const int nb = P->num_blocks;
  M->off.resize(nb + 1); M->lsize.resize(nb);
  int64_t o = 0;
  for (int b = 0; b < nb; ++b) {
    M->off[b] = o; o += P->block_size[b];
    M->lsize[b] = P->block_manifold[b] == BSNLS_MANIFOLD_QUATERNION ? 3 : P->block_size[b];
  }
  M->off[nb] = o;
  M->x.assign(P->block_values, P->block_values + o);
  int64_t total = 0;
  for (int g = 0; g < P->num_groups; ++g) total += P->groups[g].count;
  M->rb.resize(total);
  std::vector<char> seen(total, 0);
  for (int g = 0; g < P->num_groups; ++g) {
    const bsnls_factor_group& G = P->groups[g];
    const int nbk = NumBlocks(G.type);
    for (int i = 0; i < G.count; ++i) {
      const int ri = G.res_index[i];
      if (ri < 0 || ri >= total || seen[ri]) { M->error = "res_index is not a permutation"; return false; }
      seen[ri] = 1;
      ResBlock& r = M->rb[ri];
      r.type = G.type; r.idx = i; r.nb = nbk; r.nr = NumRes(G.type);
      for (int b = 0; b < nbk; ++b) r.blocks[b] = G.blocks[(int64_t)i * nbk + b];
      if (G.type == BSNLS_FACTOR_MARGINAL) { r.nb = 0; while (r.nb < nbk && r.blocks[r.nb] >= 0) ++r.nb; }   // used slots first, -1 after
    }
  }
  // Program reduction (ceres reduced_program: RemoveFixedBlocks)
  M->active.assign(nb, 0);
  for (auto& r : M->rb) {
    bool any = false;
    for (int b = 0; b < r.nb; ++b) any = any || !P->block_constant[r.blocks[b]];
    r.removed = !any;
    if (any) for (int b = 0; b < r.nb; ++b) if (!P->block_constant[r.blocks[b]]) M->active[r.blocks[b]] = 1;
  }
  M->tan.assign(nb, -1); M->epos.assign(nb, -1);
  for (int b = 0; b < nb; ++b) if (M->active[b] && P->block_group[b] == 0) {
    M->epos[b] = (int)M->eblocks.size(); M->eblocks.push_back(b); }
  for (int b = 0; b < nb; ++b) if (M->active[b] && P->block_group[b] != 0) M->fblocks.push_back(b);
  int t = 0;
  for (int b : M->eblocks) { M->tan[b] = t; t += M->lsize[b]; }
  M->ne_dim = t;
  for (int b : M->fblocks) { M->tan[b] = t; t += M->lsize[b]; }
  M->ntan = t; M->nf_dim = t - M->ne_dim;
  M->e_rows.assign(M->eblocks.size(), {});
  int64_t row = 0, jac = 0;
  M->fixed_cost = 0.0;
  for (size_t i = 0; i < M->rb.size(); ++i) {
    ResBlock& r = M->rb[i];
    r.row = row; row += r.nr;
    r.jac = jac; r.ncols = 0; r.eblock = -1;
    for (int b = 0; b < r.nb; ++b) {
      const int blk = r.blocks[b];
      if (P->block_constant[blk]) { r.jcols[b] = -1; continue; }
      r.jcols[b] = r.ncols; r.ncols += M->lsize[blk];
      if (M->epos[blk] >= 0) {
        if (r.eblock >= 0) { M->error = "group 0 is not an independent set"; return false; }
        r.eblock = M->epos[blk];
      }
    }
    jac += (int64_t)r.nr * r.ncols;
    if (r.removed) {
      double res[15]; EvalResBlockRaw(*M, r, M->x.data(), res, nullptr);
      double s = 0; for (int k = 0; k < r.nr; ++k) s += res[k] * res[k];
      int lt; double la; LossOf(*M, r, &lt, &la);
      double rho[3]; LossEvaluate(lt, la, s, rho);
      M->fixed_cost += 0.5 * rho[0];
    } else if (r.eblock >= 0) {
      M->e_rows[r.eblock].push_back((int)i);
    } else {
      M->f_rows.push_back((int)i);
    }
  }
  M->nrows = row; M->njac = jac;
  return true;
}

// ResidualBlock::Evaluate (ceres residual_block.cc; fuse restatement marginalize_variables.cpp:313-396):
// cost function -> x plus-Jacobian -> loss corrector. Jt = tangent Jacobian [nr x ncols] row-major.
static double EvalResBlock(const Model& M, const ResBlock& r, const double* x, double* res, double* Jt) {
  double jstore[15 * 32];
  double* jac[10];
  if (Jt) {
    int o = 0;
    for (int b = 0; b < r.nb; ++b) {
      const int sz = M.P->block_size[r.blocks[b]];
      jac[b] = (r.jcols[b] >= 0) ? jstore + o : nullptr;
      o += r.nr * sz;
    }
  }
  EvalResBlockRaw(M, r, x, res, Jt ? jac : nullptr);
  if (Jt) {
    for (int b = 0; b < r.nb; ++b) {
      if (r.jcols[b] < 0) continue;
      const int blk = r.blocks[b];
      const int sz = M.P->block_size[blk], ls = M.lsize[blk];
      if (M.P->block_manifold[blk] == BSNLS_MANIFOLD_QUATERNION) {
        double Pj[12]; QuatPlusJacobian(x + M.off[blk], Pj);
        for (int i = 0; i < r.nr; ++i)
          for (int k = 0; k < 3; ++k) {
            double acc = 0.0;
            for (int a = 0; a < 4; ++a) acc += jac[b][i * 4 + a] * Pj[a * 3 + k];
            Jt[i * r.ncols + r.jcols[b] + k] = acc;
          }
      } else {
        for (int i = 0; i < r.nr; ++i)
          for (int k = 0; k < ls; ++k) Jt[i * r.ncols + r.jcols[b] + k] = jac[b][i * sz + k];
      }
    }
  }
  double s = 0.0;
  for (int k = 0; k < r.nr; ++k) s += res[k] * res[k];
  int lt; double la; LossOf(M, r, &lt, &la);
  if (lt == BSNLS_LOSS_NONE) return 0.5 * s;
  double rho[3]; LossEvaluate(lt, la, s, rho);
  // ceres Corrector (corrector.cc)
  const double sqrt_rho1 = std::sqrt(rho[1]);
  double residual_scaling, alpha_sq_norm;
  if (s == 0.0 || rho[2] <= 0.0) { residual_scaling = sqrt_rho1; alpha_sq_norm = 0.0; }
  else {
    const double D = 1.0 + 2.0 * s * rho[2] / rho[1];
    const double alpha = 1.0 - std::sqrt(D);
    residual_scaling = sqrt_rho1 / (1 - alpha); alpha_sq_norm = alpha / s;
  }
  if (Jt) {
    if (alpha_sq_norm == 0.0) { for (int k = 0; k < r.nr * r.ncols; ++k) Jt[k] *= sqrt_rho1; }
    else {
      for (int c = 0; c < r.ncols; ++c) {
        double rtj = 0.0;
        for (int i = 0; i < r.nr; ++i) rtj += Jt[i * r.ncols + c] * res[i];
        for (int i = 0; i < r.nr; ++i)
          Jt[i * r.ncols + c] = sqrt_rho1 * (Jt[i * r.ncols + c] - alpha_sq_norm * res[i] * rtj);
      }
    }
  }
  for (int k = 0; k < r.nr; ++k) res[k] *= residual_scaling;
  return 0.5 * rho[0];
}

struct Workspace {
  std::vector<double> res, jac, grad, costs;
};

// Evaluator::Evaluate: cost (without fixed_cost), residuals, tangent Jacobian, gradient = J^T r.
static double Evaluate(const Model& M, const double* x, Workspace* W, bool want_jac) {
  const int64_t n = (int64_t)M.rb.size();
  W->res.resize(M.nrows); W->costs.assign(n, 0.0);
  if (want_jac) W->jac.resize(M.njac);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const ResBlock& r = M.rb[i];
    if (r.removed) {  // keep the residual vector complete for Problem::Evaluate-style output
      EvalResBlock(M, r, x, W->res.data() + r.row, nullptr);
      continue;
    }
    W->costs[i] = EvalResBlock(M, r, x, W->res.data() + r.row, want_jac ? W->jac.data() + r.jac : nullptr);
  }
  double cost = 0.0;
  for (int64_t i = 0; i < n; ++i) cost += W->costs[i];
  if (want_jac) {
    // gradient = J^T r: per-thread partial vectors over static slices of the residual blocks, summed in thread order
    // (deterministic for a given thread count; Ceres' ProgramEvaluator does the same with per-thread scratch)
    const int nth = omp_get_max_threads();
    std::vector<std::vector<double>> part(nth);
#pragma omp parallel
    {
      std::vector<double>& g = part[omp_get_thread_num()]; g.assign(M.ntan, 0.0);
#pragma omp for schedule(static)
      for (int64_t i = 0; i < n; ++i) {
        const ResBlock& r = M.rb[i];
        if (r.removed) continue;
        const double* J = W->jac.data() + r.jac; const double* rs = W->res.data() + r.row;
        for (int b = 0; b < r.nb; ++b) {
          if (r.jcols[b] < 0) continue;
          const int blk = r.blocks[b];
          for (int k = 0; k < M.lsize[blk]; ++k) {
            double acc = 0.0;
            for (int rr = 0; rr < r.nr; ++rr) acc += J[rr * r.ncols + r.jcols[b] + k] * rs[rr];
            g[M.tan[blk] + k] += acc;
          }
        }
      }
    }
    W->grad.assign(M.ntan, 0.0);
    for (int t = 0; t < nth; ++t) if (!part[t].empty()) for (int k = 0; k < M.ntan; ++k) W->grad[k] += part[t][k];
  }
  return cost;
}

// Evaluator::Plus over the reduced program's blocks
static void Plus(const Model& M, const double* x, const double* delta, double* out) {
  std::memcpy(out, x, sizeof(double) * M.off.back());
  for (int b = 0; b < M.P->num_blocks; ++b) {
    if (!M.active[b]) continue;
    const double* d = delta + M.tan[b];
    if (M.P->block_manifold[b] == BSNLS_MANIFOLD_QUATERNION) QuatPlus(x + M.off[b], d, out + M.off[b]);
    else for (int k = 0; k < M.P->block_size[b]; ++k) out[M.off[b] + k] = x[M.off[b] + k] + d[k];
  }
}
static double ActiveNorm(const Model& M, const double* a, const double* b /*nullable*/, bool inf) {
  double acc = 0.0;
  for (int blk = 0; blk < M.P->num_blocks; ++blk) {
    if (!M.active[blk]) continue;
    for (int64_t k = M.off[blk]; k < M.off[blk + 1]; ++k) {
      const double v = b ? a[k] - b[k] : a[k];
      if (inf) acc = std::max(acc, std::fabs(v)); else acc += v * v;
    }
  }
  return inf ? acc : std::sqrt(acc);
}

static void ColumnSqNorm(const Model& M, const Workspace& W, std::vector<double>* out) {
  const int nth = omp_get_max_threads();
  const int64_t n = (int64_t)M.rb.size();
  std::vector<std::vector<double>> part(nth);
#pragma omp parallel
  {
    std::vector<double>& c = part[omp_get_thread_num()]; c.assign(M.ntan, 0.0);
#pragma omp for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
      const ResBlock& r = M.rb[i];
      if (r.removed) continue;
      const double* J = W.jac.data() + r.jac;
      for (int b = 0; b < r.nb; ++b) {
        if (r.jcols[b] < 0) continue;
        const int blk = r.blocks[b];
        for (int k = 0; k < M.lsize[blk]; ++k) {
          double acc = 0.0;
          for (int rr = 0; rr < r.nr; ++rr) { const double v = J[rr * r.ncols + r.jcols[b] + k]; acc += v * v; }
          c[M.tan[blk] + k] += acc;
        }
      }
    }
  }
  out->assign(M.ntan, 0.0);
  for (int t = 0; t < nth; ++t) if (!part[t].empty()) for (int k = 0; k < M.ntan; ++k) (*out)[k] += part[t][k];
}
static void ScaleColumns(const Model& M, Workspace* W, const std::vector<double>& scale) {
  const int64_t n = (int64_t)M.rb.size();
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const ResBlock& r = M.rb[i];
    if (r.removed) continue;
    double* J = W->jac.data() + r.jac;
    for (int b = 0; b < r.nb; ++b) {
      if (r.jcols[b] < 0) continue;
      const int blk = r.blocks[b];
      for (int k = 0; k < M.lsize[blk]; ++k)
        for (int rr = 0; rr < r.nr; ++rr) J[rr * r.ncols + r.jcols[b] + k] *= scale[M.tan[blk] + k];
    }
  }
}

// Dense Cholesky (lower, row-major n x n), returns false when not positive definite.
static bool CholeskyLower(std::vector<double>& A, int n) {
  for (int i = 0; i < n; ++i) {
    double* Ai = &A[(size_t)i * n];
    for (int j = 0; j <= i; ++j) {
      const double* Aj = &A[(size_t)j * n];
      double s = Ai[j];
      for (int k = 0; k < j; ++k) s -= Ai[k] * Aj[k];
      if (i == j) { if (!(s > 0.0) || !std::isfinite(s)) return false; Ai[i] = std::sqrt(s); }
      else Ai[j] = s / Aj[j];
    }
  }
  return true;
}
static void CholeskySolve(const std::vector<double>& L, int n, double* b) {
  for (int i = 0; i < n; ++i) {
    double s = b[i]; const double* Li = &L[(size_t)i * n];
    for (int k = 0; k < i; ++k) s -= Li[k] * b[k];
    b[i] = s / Li[i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = b[i];
    for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * b[k];
    b[i] = s / L[(size_t)i * n + i];
  }
}
// inverse of a small SPD matrix via Cholesky (ceres InvertPSDMatrix, fixed-size path)
static bool InvertSPD(const double* A, int n, double* inv) {
  double L[9];
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = A[i * n + j];
      for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
      if (i == j) { if (!(s > 0.0)) return false; L[i * n + i] = std::sqrt(s); } else L[i * n + j] = s / L[j * n + j];
    }
  for (int c = 0; c < n; ++c) {
    double y[3];
    for (int i = 0; i < n; ++i) { double s = (i == c) ? 1.0 : 0.0; for (int k = 0; k < i; ++k) s -= L[i * n + k] * y[k]; y[i] = s / L[i * n + i]; }
    for (int i = n - 1; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < n; ++k) s -= L[k * n + i] * y[k]; y[i] = s / L[i * n + i]; }
    for (int i = 0; i < n; ++i) inv[i * n + c] = y[i];
  }
  return true;
}

// Linear-algebra mode of the reduced system: 0 = automatic (dense up to kDenseLimit unknowns, block-sparse above),
// 1 = always dense, 2 = always block-sparse. Set through oracle_set_linear_mode (tests cross-check the two).
static int g_linear_mode = 0;
static const int kDenseLimit = 2400;
static bool UseSparse(const Model& M) { return g_linear_mode == 2 || (g_linear_mode == 0 && M.nf_dim > kDenseLimit); }

// SchurEliminator::Eliminate (schur_eliminator_impl.h): reduced system over f-blocks in program order, stored dense
// (small windows) or block-sparse on the structural pattern (oracle_sparse.h).
// J is whatever W holds (Jacobi-scaled in the LM loop), D = LM diagonal (per tangent column, e then f).
struct Reduced {
  bool sparse = false;
  std::vector<double> S, rhs;     // dense: nf x nf row-major (full symmetric); sparse: block values of `pat`
  BlockPattern pat; SparseCholesky chol; bool have_pattern = false;
  std::vector<double> ete_inv;    // per e-block ls x ls
  std::vector<double> ge;         // E^T r per e-block
};
static void BuildPattern(const Model& M, Reduced* R, bool analyse = true) {
  const int nfb = (int)M.fblocks.size(), ne0 = M.ne_dim;
  std::vector<int> fpos(M.P->num_blocks, -1), ls(nfb);
  for (int k = 0; k < nfb; ++k) { fpos[M.fblocks[k]] = k; ls[k] = M.lsize[M.fblocks[k]]; }
  LazyAdj A(nfb);
  std::vector<int> cl;
  for (size_t e = 0; e < M.eblocks.size(); ++e) {
    cl.clear();
    for (int ri : M.e_rows[e]) { const ResBlock& r = M.rb[ri]; for (int b = 0; b < r.nb; ++b) if (r.jcols[b] >= 0 && fpos[r.blocks[b]] >= 0) cl.push_back(fpos[r.blocks[b]]); }
    std::sort(cl.begin(), cl.end()); cl.erase(std::unique(cl.begin(), cl.end()), cl.end());
    A.AddClique(cl);
  }
  for (int ri : M.f_rows) {
    const ResBlock& r = M.rb[ri]; cl.clear();
    for (int b = 0; b < r.nb; ++b) if (r.jcols[b] >= 0 && fpos[r.blocks[b]] >= 0) cl.push_back(fpos[r.blocks[b]]);
    A.AddClique(cl);
  }
  FinishPattern(&A, ls, &R->pat);
  (void)ne0;
  if (analyse) R->chol.Analyse(R->pat);   // ordering + symbolic factorisation: only needed when the system is solved
  R->have_pattern = true;
}
struct DenseAcc {
  double* S; int nf;
  inline void Add(int r, int c, double v) { S[(size_t)r * nf + c] += v; }   // r <= c
  inline bool SameBlock(int a, int b) const { return a / 3 == b / 3; }     // only groups the loops; any grouping is valid
  void MirrorUpperToLower() {
#pragma omp parallel for schedule(static)
    for (int r = 0; r < nf; ++r) for (int c = r + 1; c < nf; ++c) S[(size_t)c * nf + r] = S[(size_t)r * nf + c];
  }
};
struct SparseAcc {
  double* V; const BlockPattern* P; int lr = -1, lc = -1; int64_t lo = 0; int ljs = 0;
  inline void Add(int r, int c, double v) {
    const int br = P->blk_of[r], bc = P->blk_of[c];
    if (br != lr || bc != lc) { lr = br; lc = bc; lo = P->voff[P->Slot(br, bc)]; ljs = P->ls[bc]; }
    V[lo + (int64_t)(r - P->off[br]) * ljs + (c - P->off[bc])] += v;   // r <= c
  }
  inline bool SameBlock(int a, int b) const { return P->blk_of[a] == P->blk_of[b]; }
  void MirrorUpperToLower() {
#pragma omp parallel for schedule(static)
    for (int bi = 0; bi < P->nfb; ++bi)
      for (int s = P->rowptr[bi]; s < P->rowptr[bi + 1]; ++s) {
        const int bj = P->col[s]; if (bj < bi) continue;
        const int li = P->ls[bi], lj = P->ls[bj];
        double* U = V + P->voff[s];
        if (bj == bi) { for (int a = 0; a < li; ++a) for (int b = a + 1; b < li; ++b) U[b * li + a] = U[a * li + b]; continue; }
        double* L = V + P->voff[P->Slot(bj, bi)];
        for (int a = 0; a < li; ++a) for (int b = 0; b < lj; ++b) L[b * li + a] = U[a * lj + b];
      }
  }
};
static inline void Init(DenseAcc& a, double* S, int nf, const BlockPattern*) { a.S = S; a.nf = nf; }
static inline void Init(SparseAcc& a, double* S, int, const BlockPattern* P) { a.V = S; a.P = P; }
template <class Acc>
static bool BuildReducedT(const Model& M, const Workspace& W, const double* D, Reduced* R, size_t nS) {
  const int nf = M.nf_dim, ne0 = M.ne_dim;
  const int nthreads = omp_get_max_threads();
  std::vector<std::vector<double>> Sloc(nthreads), rloc(nthreads);
  R->ete_inv.assign((size_t)M.eblocks.size() * 9, 0.0); R->ge.assign((size_t)M.eblocks.size() * 3, 0.0);
  bool ok = true;
  auto make_acc = [&](double* S) { Acc a{}; Init(a, S, nf, &R->pat); return a; };
#pragma omp parallel
  {
    const int tid = omp_get_thread_num();
    std::vector<double>& S = Sloc[tid]; std::vector<double>& rhs = rloc[tid];
    S.assign(nS, 0.0); rhs.assign(nf, 0.0);
    Acc acc = make_acc(S.data());
#pragma omp for schedule(static)
    for (int64_t e = 0; e < (int64_t)M.eblocks.size(); ++e) {
      const int eb = M.eblocks[e], le = M.lsize[eb];
      double ete[9] = {0}, g[3] = {0};
      if (D) for (int k = 0; k < le; ++k) ete[k * le + k] = D[M.tan[eb] + k] * D[M.tan[eb] + k];
      // pass 1: ete, g
      for (int ri : M.e_rows[e]) {
        const ResBlock& r = M.rb[ri];
        const double* J = W.jac.data() + r.jac; const double* rs = W.res.data() + r.row;
        int ec = -1; for (int b = 0; b < r.nb; ++b) if (r.jcols[b] >= 0 && r.blocks[b] == eb) ec = r.jcols[b];
        for (int a = 0; a < le; ++a) {
          for (int c = 0; c < le; ++c) { double acc2 = 0; for (int rr = 0; rr < r.nr; ++rr) acc2 += J[rr * r.ncols + ec + a] * J[rr * r.ncols + ec + c]; ete[a * le + c] += acc2; }
          double acc2 = 0; for (int rr = 0; rr < r.nr; ++rr) acc2 += J[rr * r.ncols + ec + a] * rs[rr]; g[a] += acc2;
        }
      }
      double inv[9];
      if (!InvertSPD(ete, le, inv)) { ok = false; continue; }
      for (int k = 0; k < le * le; ++k) R->ete_inv[e * 9 + k] = inv[k];
      for (int k = 0; k < le; ++k) R->ge[e * 3 + k] = g[k];
      double invg[3]; for (int a = 0; a < le; ++a) { double acc2 = 0; for (int c = 0; c < le; ++c) acc2 += inv[a * le + c] * g[c]; invg[a] = acc2; }
      // pass 2: F^T F, F^T r, and collect F^T E per (row block, f column)
      struct Col { int gcol; double fte[3]; };
      std::vector<Col> cols;
      for (int ri : M.e_rows[e]) {
        const ResBlock& r = M.rb[ri];
        const double* J = W.jac.data() + r.jac; const double* rs = W.res.data() + r.row;
        int ec = -1; for (int b = 0; b < r.nb; ++b) if (r.jcols[b] >= 0 && r.blocks[b] == eb) ec = r.jcols[b];
        for (int b = 0; b < r.nb; ++b) {
          if (r.jcols[b] < 0 || r.blocks[b] == eb) continue;
          const int blk = r.blocks[b], gc0 = M.tan[blk] - ne0;
          for (int k = 0; k < M.lsize[blk]; ++k) {
            const int lc = r.jcols[b] + k;
            double acc2 = 0; for (int rr = 0; rr < r.nr; ++rr) acc2 += J[rr * r.ncols + lc] * rs[rr];
            rhs[gc0 + k] += acc2;
            // F^T F against every f column of the same residual block (upper triangle; mirrored at the end, as Ceres'
            // SchurEliminator only forms the block pairs i <= j)
            for (int b2 = 0; b2 < r.nb; ++b2) {
              if (r.jcols[b2] < 0 || r.blocks[b2] == eb) continue;
              const int blk2 = r.blocks[b2], gc2 = M.tan[blk2] - ne0;
              for (int k2 = 0; k2 < M.lsize[blk2]; ++k2) {
                if (gc2 + k2 < gc0 + k) continue;
                double a2 = 0; for (int rr = 0; rr < r.nr; ++rr) a2 += J[rr * r.ncols + lc] * J[rr * r.ncols + r.jcols[b2] + k2];
                acc.Add(gc0 + k, gc2 + k2, a2);
              }
            }
            // F^T E accumulated per global column (a camera may be hit by several rows of the chunk)
            Col* cptr = nullptr;
            for (auto& c : cols) if (c.gcol == gc0 + k) { cptr = &c; break; }
            if (!cptr) { cols.push_back({gc0 + k, {0, 0, 0}}); cptr = &cols.back(); }
            for (int a = 0; a < le; ++a) { double a3 = 0; for (int rr = 0; rr < r.nr; ++rr) a3 += J[rr * r.ncols + lc] * J[rr * r.ncols + ec + a]; cptr->fte[a] += a3; }
          }
        }
      }
      // S -= (F^T E) inv (E^T F); rhs -= (F^T E) inv g.  Columns sorted so that a block's columns are adjacent (one block
      // lookup per block pair) and only the upper triangle is formed.
      std::sort(cols.begin(), cols.end(), [](const Col& a, const Col& b) { return a.gcol < b.gcol; });
      std::vector<double> tt(cols.size() * 3);
      for (size_t i = 0; i < cols.size(); ++i) {
        const Col& ci = cols[i];
        for (int a = 0; a < le; ++a) { double acc2 = 0; for (int c = 0; c < le; ++c) acc2 += ci.fte[c] * inv[c * le + a]; tt[3 * i + a] = acc2; }
        double acc2 = 0; for (int a = 0; a < le; ++a) acc2 += ci.fte[a] * invg[a];
        rhs[ci.gcol] -= acc2;
      }
      for (size_t i0 = 0; i0 < cols.size();) {
        size_t i1 = i0 + 1; while (i1 < cols.size() && acc.SameBlock(cols[i1].gcol, cols[i0].gcol)) ++i1;
        for (size_t j0 = i0; j0 < cols.size();) {
          size_t j1 = j0 + 1; while (j1 < cols.size() && acc.SameBlock(cols[j1].gcol, cols[j0].gcol)) ++j1;
          for (size_t i = i0; i < i1; ++i)
            for (size_t j = std::max(j0, i); j < j1; ++j) {
              double a2 = 0; for (int a = 0; a < le; ++a) a2 += tt[3 * i + a] * cols[j].fte[a];
              acc.Add(cols[i].gcol, cols[j].gcol, -a2);
            }
          j0 = j1;
        }
        i0 = i1;
      }
    }
  }
  if (!ok) return false;
  R->S.assign(nS, 0.0); R->rhs.assign(nf, 0.0);
  for (int t = 0; t < nthreads; ++t) {
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < (int64_t)nS; ++k) R->S[k] += Sloc[t][k];
    for (int k = 0; k < nf; ++k) R->rhs[k] += rloc[t][k];
    std::vector<double>().swap(Sloc[t]);
  }
  Acc acc = make_acc(R->S.data());
  // rows without an e-block (NoEBlockRowsUpdate) and D_f^2
  for (int ri : M.f_rows) {
    const ResBlock& r = M.rb[ri];
    const double* J = W.jac.data() + r.jac; const double* rs = W.res.data() + r.row;
    for (int b = 0; b < r.nb; ++b) {
      if (r.jcols[b] < 0) continue;
      const int blk = r.blocks[b], gc0 = M.tan[blk] - ne0;
      for (int k = 0; k < M.lsize[blk]; ++k) {
        const int lc = r.jcols[b] + k;
        double acc2 = 0; for (int rr = 0; rr < r.nr; ++rr) acc2 += J[rr * r.ncols + lc] * rs[rr];
        R->rhs[gc0 + k] += acc2;
        for (int b2 = 0; b2 < r.nb; ++b2) {
          if (r.jcols[b2] < 0) continue;
          const int blk2 = r.blocks[b2], gc2 = M.tan[blk2] - ne0;
          for (int k2 = 0; k2 < M.lsize[blk2]; ++k2) {
            if (gc2 + k2 < gc0 + k) continue;
            double a2 = 0; for (int rr = 0; rr < r.nr; ++rr) a2 += J[rr * r.ncols + lc] * J[rr * r.ncols + r.jcols[b2] + k2];
            acc.Add(gc0 + k, gc2 + k2, a2);
          }
        }
      }
    }
  }
  if (D) for (int k = 0; k < nf; ++k) acc.Add(k, k, D[ne0 + k] * D[ne0 + k]);
  acc.MirrorUpperToLower();
  return true;
}
static bool BuildReduced(const Model& M, const Workspace& W, const double* D, Reduced* R) {
  R->sparse = UseSparse(M);
  if (!R->sparse) return BuildReducedT<DenseAcc>(M, W, D, R, (size_t)M.nf_dim * M.nf_dim);
  if (!R->have_pattern) BuildPattern(M, R);
  return BuildReducedT<SparseAcc>(M, W, D, R, (size_t)R->pat.nvals);
}
// Solve S y = rhs by Cholesky (dense LL^T, or the sparse factorisation on the block pattern); false if S is not SPD.
static bool SolveReduced(const Model& M, Reduced* R, std::vector<double>* yf) {
  *yf = R->rhs;
  if (R->sparse) { if (!R->chol.analysed) R->chol.Analyse(R->pat); if (!R->chol.Factor(R->S)) return false; R->chol.Solve(yf->data()); return true; }
  std::vector<double> L = R->S;
  if (!CholeskyLower(L, M.nf_dim)) return false;
  CholeskySolve(L, M.nf_dim, yf->data());
  return true;
}

// SchurEliminator::BackSubstitute: y_e = ete^-1 (E^T r - E^T F y_f)
static void BackSubstitute(const Model& M, const Workspace& W, const Reduced& R, const double* yf, double* y) {
  const int ne0 = M.ne_dim;
  for (int k = 0; k < M.nf_dim; ++k) y[ne0 + k] = yf[k];
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < (int64_t)M.eblocks.size(); ++e) {
    const int eb = M.eblocks[e], le = M.lsize[eb];
    double g[3]; for (int k = 0; k < le; ++k) g[k] = R.ge[e * 3 + k];
    for (int ri : M.e_rows[e]) {
      const ResBlock& r = M.rb[ri];
      const double* J = W.jac.data() + r.jac;
      int ec = -1; for (int b = 0; b < r.nb; ++b) if (r.jcols[b] >= 0 && r.blocks[b] == eb) ec = r.jcols[b];
      double Fy[15] = {0};
      for (int b = 0; b < r.nb; ++b) {
        if (r.jcols[b] < 0 || r.blocks[b] == eb) continue;
        const int blk = r.blocks[b];
        for (int k = 0; k < M.lsize[blk]; ++k)
          for (int rr = 0; rr < r.nr; ++rr) Fy[rr] += J[rr * r.ncols + r.jcols[b] + k] * yf[M.tan[blk] - ne0 + k];
      }
      for (int a = 0; a < le; ++a) { double acc = 0; for (int rr = 0; rr < r.nr; ++rr) acc += J[rr * r.ncols + ec + a] * Fy[rr]; g[a] -= acc; }
    }
    for (int a = 0; a < le; ++a) { double acc = 0; for (int c = 0; c < le; ++c) acc += R.ete_inv[e * 9 + a * le + c] * g[c]; y[M.tan[eb] + a] = acc; }
  }
}

static void SetMessage(bsnls_summary* s, const char* m) { std::snprintf(s->message, sizeof(s->message), "%s", m); }

// ceres TrustRegionMinimizer::Minimize with LevenbergMarquardtStrategy (monotonic steps, no inner
// iterations, no bounds — none of which the reference enables, ceres_options.cpp:75-200).
// trace (optional): the parameter vector at the start of every attempted step, one after the other (iteration k's
// linearisation point is entry k-1), for the teacher-forced per-step parity test.
static int Solve(const bsnls_problem* P, const bsnls_options* opt, bsnls_summary* sum, bsnls_iteration* its,
                 int max_its, double* x_out, std::vector<double>* trace = nullptr) {
  const auto t0 = std::chrono::steady_clock::now();
  Model M;
  std::memset(sum, 0, sizeof(*sum));
  if (!BuildModel(P, &M)) { SetMessage(sum, M.error.c_str()); sum->termination_type = BSNLS_FAILURE; return BSNLS_ERR_INVALID; }
  Workspace W;
  std::vector<bsnls_iteration> iters;
  std::vector<double> x = M.x, cand(x.size()), scale, diag, lmD(M.ntan), step(M.ntan), delta(M.ntan), tmp(x.size());
  sum->fixed_cost = M.fixed_cost;
  sum->num_e_blocks = (int)M.eblocks.size(); sum->num_f_blocks = (int)M.fblocks.size();
  sum->reduced_system_dim = M.nf_dim; sum->num_effective_parameters_reduced = M.ntan;
  sum->num_parameter_blocks_reduced = sum->num_e_blocks + sum->num_f_blocks;
  for (const auto& r : M.rb) if (!r.removed) { sum->num_residual_blocks_reduced++; sum->num_residuals_reduced += r.nr; }
  for (int b = 0; b < P->num_blocks; ++b) if (M.active[b]) sum->num_parameters_reduced += P->block_size[b];
  sum->termination_type = BSNLS_NO_CONVERGENCE;

  double radius = opt->initial_trust_region_radius, decrease_factor = 2.0;
  bool reuse_diagonal = false;
  int num_invalid = 0;
  // IterationZero
  bsnls_iteration it; std::memset(&it, 0, sizeof(it));
  double x_cost = Evaluate(M, x.data(), &W, true);
  double x_norm = ActiveNorm(M, x.data(), nullptr, false);
  auto gradient_norms = [&](bsnls_iteration* s) {
    std::vector<double> ng(M.ntan); for (int k = 0; k < M.ntan; ++k) ng[k] = -W.grad[k];
    Plus(M, x.data(), ng.data(), tmp.data());
    s->gradient_max_norm = ActiveNorm(M, x.data(), tmp.data(), true);
    s->gradient_norm = ActiveNorm(M, x.data(), tmp.data(), false);
  };
  if (opt->jacobi_scaling) {
    ColumnSqNorm(M, W, &scale);
    for (auto& s : scale) s = 1.0 / (1.0 + std::sqrt(s));
    ScaleColumns(M, &W, scale);
  } else scale.assign(M.ntan, 1.0);
  it.cost = x_cost + M.fixed_cost; gradient_norms(&it);
  it.step_is_valid = 1; it.step_is_successful = 1;
  sum->initial_cost = it.cost;
  Reduced R;
  const char* msg = "";
  for (;;) {
    // FinalizeIterationAndCheckIfMinimizerCanContinue
    if (it.step_is_successful) sum->num_successful_steps++;
    it.trust_region_radius = radius;
    iters.push_back(it);
    const double elapsed = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (elapsed >= opt->max_solver_time_in_seconds) { sum->termination_type = BSNLS_NO_CONVERGENCE; msg = "Maximum solver time reached."; break; }
    if (it.iteration >= opt->max_num_iterations) { sum->termination_type = BSNLS_NO_CONVERGENCE; msg = "Maximum number of iterations reached."; break; }
    if (it.step_is_successful && it.gradient_max_norm <= opt->gradient_tolerance) { sum->termination_type = BSNLS_CONVERGENCE; msg = "Gradient tolerance reached."; break; }
    if (radius < opt->min_trust_region_radius) { sum->termination_type = BSNLS_CONVERGENCE; msg = "Minimum trust region radius reached."; break; }
    const bsnls_iteration prev = it;
    std::memset(&it, 0, sizeof(it)); it.iteration = prev.iteration + 1;
    if (trace) trace->insert(trace->end(), x.begin(), x.end());
    // ComputeTrustRegionStep — LevenbergMarquardtStrategy::ComputeStep
    if (!reuse_diagonal) {
      ColumnSqNorm(M, W, &diag);
      for (auto& d : diag) d = std::min(std::max(d, opt->min_lm_diagonal), opt->max_lm_diagonal);
    }
    for (int k = 0; k < M.ntan; ++k) lmD[k] = std::sqrt(diag[k] / radius);
    bool solved = BuildReduced(M, W, lmD.data(), &R);
    std::vector<double> yf;
    if (solved) solved = SolveReduced(M, &R, &yf);
    if (solved) {
      BackSubstitute(M, W, R, yf.data(), step.data());
      for (int k = 0; k < M.ntan; ++k) { if (!std::isfinite(step[k])) solved = false; step[k] = -step[k]; }
    }
    reuse_diagonal = true;
    double model_cost_change = 0.0;
    if (solved) {
      // model_cost_change = -(J step)^T (f + J step / 2)
      for (const ResBlock& r : M.rb) {
        if (r.removed) continue;
        const double* J = W.jac.data() + r.jac; const double* rs = W.res.data() + r.row;
        for (int rr = 0; rr < r.nr; ++rr) {
          double m = 0.0;
          for (int b = 0; b < r.nb; ++b) {
            if (r.jcols[b] < 0) continue;
            const int blk = r.blocks[b];
            for (int k = 0; k < M.lsize[blk]; ++k) m += J[rr * r.ncols + r.jcols[b] + k] * step[M.tan[blk] + k];
          }
          model_cost_change -= m * (rs[rr] + m / 2.0);
        }
      }
      it.step_is_valid = model_cost_change > 0.0;
    }
    if (!it.step_is_valid) {
      // HandleInvalidStep
      if (++num_invalid >= opt->max_num_consecutive_invalid_steps) {
        sum->termination_type = BSNLS_FAILURE; msg = "Number of consecutive invalid steps more than Solver::Options::max_num_consecutive_invalid_steps.";
        break;
      }
      radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
      it.cost = prev.cost; it.cost_change = 0.0; it.gradient_max_norm = prev.gradient_max_norm;
      it.gradient_norm = prev.gradient_norm; it.step_norm = 0.0; it.relative_decrease = 0.0;
      sum->num_unsuccessful_steps++;
      continue;
    }
    num_invalid = 0;
    for (int k = 0; k < M.ntan; ++k) delta[k] = step[k] * scale[k];
    // ComputeCandidatePointAndEvaluateCost
    Plus(M, x.data(), delta.data(), cand.data());
    Workspace Wc;
    const double cand_cost = Evaluate(M, cand.data(), &Wc, false);
    // ParameterToleranceReached
    it.step_norm = ActiveNorm(M, x.data(), cand.data(), false);
    it.gradient_max_norm = prev.gradient_max_norm; it.gradient_norm = prev.gradient_norm;
    if (it.step_norm <= opt->parameter_tolerance * (x_norm + opt->parameter_tolerance)) {
      // returns from Minimize without finalising: this iteration is not recorded, x stays put
      sum->termination_type = BSNLS_CONVERGENCE; msg = "Parameter tolerance reached."; break;
    }
    // FunctionToleranceReached
    it.cost_change = x_cost - cand_cost;
    if (std::fabs(it.cost_change) <= opt->function_tolerance * x_cost) {
      sum->termination_type = BSNLS_CONVERGENCE; msg = "Function tolerance reached."; break;
    }
    it.relative_decrease = (x_cost - cand_cost) / model_cost_change;
    if (it.relative_decrease > opt->min_relative_decrease) {
      // HandleSuccessfulStep
      x = cand; x_norm = ActiveNorm(M, x.data(), nullptr, false);
      x_cost = Evaluate(M, x.data(), &W, true);
      it.cost = x_cost + M.fixed_cost;
      gradient_norms(&it);
      if (opt->jacobi_scaling) ScaleColumns(M, &W, scale);
      it.step_is_successful = 1;
      radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * it.relative_decrease - 1.0, 3));
      radius = std::min(opt->max_trust_region_radius, radius);
      decrease_factor = 2.0; reuse_diagonal = false;
    } else {
      // HandleUnsuccessfulStep
      it.step_is_successful = 0; sum->num_unsuccessful_steps++;
      radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
      it.cost = cand_cost + M.fixed_cost;
    }
  }
  SetMessage(sum, msg);
  sum->num_iterations = (int)iters.size();
  sum->final_cost = sum->initial_cost;
  for (const auto& i : iters) sum->final_cost = std::min(sum->final_cost, i.cost);  // solver.cc SetSummaryFinalCost
  for (int k = 0; k < (int)iters.size() && k < max_its; ++k) if (its) its[k] = iters[k];
  if (x_out) std::memcpy(x_out, x.data(), sizeof(double) * x.size());
  sum->linear_solver_used = BSNLS_LINEAR_DENSE_SCHUR;
  sum->total_time_in_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return BSNLS_OK;
}

}  // namespace orc

extern "C" {

int oracle_solve(const bsnls_problem* P, const bsnls_options* opt, bsnls_summary* sum, bsnls_iteration* its,
                 int32_t max_its, double* x_out) {
  return orc::Solve(P, opt, sum, its, max_its, x_out);
}

// Problem::Evaluate (hash_graph.cpp:446-455): cost incl. constant residual blocks, residuals of every residual
// block in program order (loss-corrected), tangent gradient over the non-constant blocks in PROGRAM order.
int oracle_evaluate(const bsnls_problem* P, double* cost, double* residuals, int64_t* nres, double* gradient,
                    int64_t* ngrad) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return BSNLS_ERR_INVALID;
  orc::Workspace W;
  const double c = orc::Evaluate(M, M.x.data(), &W, true);
  if (cost) *cost = c + M.fixed_cost;
  if (nres) *nres = M.nrows;
  if (residuals) std::memcpy(residuals, W.res.data(), sizeof(double) * M.nrows);
  int64_t n = 0;
  for (int b = 0; b < P->num_blocks; ++b) {
    if (!M.active[b]) continue;
    if (gradient) for (int k = 0; k < M.lsize[b]; ++k) gradient[n + k] = W.grad[M.tan[b] + k];
    n += M.lsize[b];
  }
  if (ngrad) *ngrad = n;
  return BSNLS_OK;
}

// One residual block: raw CostFunction::Evaluate output (ambient Jacobians, row-major, concatenated by block)
int oracle_cost_function(const bsnls_problem* P, int32_t type, int32_t index, double* residual, double* jac) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return BSNLS_ERR_INVALID;
  for (const auto& r : M.rb) {
    if (r.type != type || r.idx != index) continue;
    double* jp[10]; int o = 0;
    for (int b = 0; b < r.nb; ++b) { jp[b] = jac + o; o += r.nr * P->block_size[r.blocks[b]]; }
    orc::EvalResBlockRaw(M, r, M.x.data(), residual, jac ? jp : nullptr);
    return BSNLS_OK;
  }
  return BSNLS_ERR_INVALID;
}
// One residual block linearised like fuse_constraints::detail::linearize (marginalize_variables.cpp:281-399):
// tangent Jacobian [nr x ncols] row-major over its non-constant blocks + loss-corrected residual.
int oracle_linearize(const bsnls_problem* P, int32_t type, int32_t index, double* residual, double* jac,
                     int32_t* ncols) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return BSNLS_ERR_INVALID;
  for (const auto& r : M.rb) {
    if (r.type != type || r.idx != index) continue;
    orc::EvalResBlock(M, r, M.x.data(), residual, jac);
    if (ncols) *ncols = r.ncols;
    return BSNLS_OK;
  }
  return BSNLS_ERR_INVALID;
}
// Dense reduced system at the current values for a given radius, in UNSCALED coordinates:
// (J^T J + D^2) with D^2_j = clamp(c_j s_j^2)/(radius s_j^2), s = Jacobi scale, after landmark elimination.
// S: nf x nf row-major full symmetric; rhs: nf.  (Compared against bsnls_get_reduced_system.)
int oracle_reduced_system(const bsnls_problem* P, const bsnls_options* opt, double radius, double* S,
                          double* rhs, int32_t* dim) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return BSNLS_ERR_INVALID;
  if (dim) *dim = M.nf_dim;
  if (!S) return BSNLS_OK;
  orc::Workspace W; orc::Evaluate(M, M.x.data(), &W, true);
  std::vector<double> c, D(M.ntan);
  orc::ColumnSqNorm(M, W, &c);
  for (int k = 0; k < M.ntan; ++k) {
    const double s = opt->jacobi_scaling ? 1.0 / (1.0 + std::sqrt(c[k])) : 1.0;
    const double d = std::min(std::max(c[k] * s * s, opt->min_lm_diagonal), opt->max_lm_diagonal);
    D[k] = std::sqrt(d / radius) / s;
  }
  orc::Reduced R;
  const int saved = orc::g_linear_mode; orc::g_linear_mode = 1;   // this entry point returns the dense matrix
  const bool ok = orc::BuildReduced(M, W, D.data(), &R);
  orc::g_linear_mode = saved;
  if (!ok) return BSNLS_ERR_INVALID;
  std::memcpy(S, R.S.data(), sizeof(double) * R.S.size());
  std::memcpy(rhs, R.rhs.data(), sizeof(double) * R.rhs.size());
  return BSNLS_OK;
}
// Same system as oracle_reduced_system for graphs whose dense matrix does not fit: the lower-triangular 3x3 (in general
// ls_r x ls_c) blocks of S on its structural pattern, sorted by (row, col) in f-block positions of the program order,
// values row-major, 9 doubles reserved per block. Call with vals == NULL to get the count.
int oracle_reduced_blocks(const bsnls_problem* P, const bsnls_options* opt, double radius, int32_t* rows, int32_t* cols,
                          double* vals, int64_t* count, double* rhs) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return BSNLS_ERR_INVALID;
  const int saved = orc::g_linear_mode; orc::g_linear_mode = 2;
  orc::Reduced R;
  if (!vals) {
    orc::BuildPattern(M, &R, false);
    int64_t n = 0;
    for (int i = 0; i < R.pat.nfb; ++i) for (int s = R.pat.rowptr[i]; s < R.pat.rowptr[i + 1]; ++s) n += R.pat.col[s] <= i;
    *count = n; orc::g_linear_mode = saved; return BSNLS_OK;
  }
  orc::Workspace W; orc::Evaluate(M, M.x.data(), &W, true);
  std::vector<double> c, D(M.ntan);
  orc::ColumnSqNorm(M, W, &c);
  for (int k = 0; k < M.ntan; ++k) {
    const double s = opt->jacobi_scaling ? 1.0 / (1.0 + std::sqrt(c[k])) : 1.0;
    const double d = std::min(std::max(c[k] * s * s, opt->min_lm_diagonal), opt->max_lm_diagonal);
    D[k] = std::sqrt(d / radius) / s;
  }
  const bool ok = orc::BuildReduced(M, W, D.data(), &R);
  orc::g_linear_mode = saved;
  if (!ok) return BSNLS_ERR_INVALID;
  int64_t n = 0;
  for (int i = 0; i < R.pat.nfb; ++i)
    for (int s = R.pat.rowptr[i]; s < R.pat.rowptr[i + 1]; ++s) {
      const int j = R.pat.col[s]; if (j > i) continue;
      if (n >= *count) return BSNLS_ERR_INVALID;
      rows[n] = i; cols[n] = j;
      for (int k = 0; k < 9; ++k) vals[9 * n + k] = 0.0;
      for (int a = 0; a < R.pat.ls[i]; ++a) for (int b = 0; b < R.pat.ls[j]; ++b) vals[9 * n + 3 * a + b] = R.S[R.pat.voff[s] + (int64_t)a * R.pat.ls[j] + b];
      ++n;
    }
  *count = n;
  if (rhs) std::memcpy(rhs, R.rhs.data(), sizeof(double) * R.rhs.size());
  return BSNLS_OK;
}
// LossFunction::Evaluate of one loss class: rho(s), rho'(s), rho''(s)
void oracle_loss(int32_t type, double a, double s, double* rho) { orc::LossEvaluate(type, a, s, rho); }
void oracle_set_linear_mode(int32_t mode) { orc::g_linear_mode = mode; }
// oracle_solve + the linearisation point of every attempted step (num_values doubles each, at most max_trace points)
int oracle_solve_trace(const bsnls_problem* P, const bsnls_options* opt, bsnls_summary* sum, bsnls_iteration* its,
                       int32_t max_its, double* x_out, double* x_trace, int32_t max_trace, int32_t* num_trace) {
  std::vector<double> tr;
  const int rc = orc::Solve(P, opt, sum, its, max_its, x_out, &tr);
  int64_t nv = 0; for (int b = 0; b < P->num_blocks; ++b) nv += P->block_size[b];
  const int n = nv ? (int)(tr.size() / nv) : 0;
  if (num_trace) *num_trace = n;
  if (x_trace) std::memcpy(x_trace, tr.data(), sizeof(double) * nv * std::min(n, max_trace));
  return rc;
}
int oracle_get_ordering(const bsnls_problem* P, int32_t* e_blocks, int32_t* ne, int32_t* f_blocks, int32_t* nf) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return BSNLS_ERR_INVALID;
  *ne = (int)M.eblocks.size(); *nf = (int)M.fblocks.size();
  if (e_blocks) std::memcpy(e_blocks, M.eblocks.data(), sizeof(int) * M.eblocks.size());
  if (f_blocks) std::memcpy(f_blocks, M.fblocks.data(), sizeof(int) * M.fblocks.size());
  return BSNLS_OK;
}
// Manifold helpers (orientation_3d_stamped.h:91-142)
void oracle_quat_plus(const double* x, const double* d, double* out) { orc::QuatPlus(x, d, out); }
void oracle_quat_minus(const double* x1, const double* x2, double* out) { orc::QuatMinus(x1, x2, out); }
void oracle_quat_plus_jacobian(const double* x, double* J) { orc::QuatPlusJacobian(x, J); }
void oracle_point_transformation(const double* q, const double* t, const double* P, double* out) {
  orc::PointTransformation(q, t, P, out);
}

// CPU baseline: `reps` passes of residual+Jacobian evaluation followed by Schur elimination into the dense
// reduced system (the work bsnls_run_eval + bsnls_run_schur do), OpenMP over all host threads.
// Returns seconds per pass (best of reps).
double oracle_time_eval_schur(const bsnls_problem* P, const bsnls_options* opt, double radius, int32_t reps,
                              int32_t* threads) {
  orc::Model M; if (!orc::BuildModel(P, &M)) return -1.0;
  if (threads) *threads = omp_get_max_threads();
  orc::Workspace W; std::vector<double> c, D(M.ntan); orc::Reduced R;
  if (orc::UseSparse(M)) orc::BuildPattern(M, &R, false);   // structure of S: fixed per graph (Ceres builds it once per Solve), not part of a pass
  double best = 1e300;
  for (int i = 0; i < reps; ++i) {
    const auto t0 = std::chrono::steady_clock::now();
    orc::Evaluate(M, M.x.data(), &W, true);
    const auto t1 = std::chrono::steady_clock::now();
    orc::ColumnSqNorm(M, W, &c);
    for (int k = 0; k < M.ntan; ++k) {
      const double s = opt->jacobi_scaling ? 1.0 / (1.0 + std::sqrt(c[k])) : 1.0;
      D[k] = std::sqrt(std::min(std::max(c[k] * s * s, opt->min_lm_diagonal), opt->max_lm_diagonal) / radius) / s;
    }
    const auto t2 = std::chrono::steady_clock::now();
    orc::BuildReduced(M, W, D.data(), &R);
    const auto t3 = std::chrono::steady_clock::now();
    if (std::getenv("ORACLE_PROFILE")) std::fprintf(stderr, "[oracle] evaluate %.3f s, column norms %.3f s, Schur elimination %.3f s\n", std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(t2 - t1).count(), std::chrono::duration<double>(t3 - t2).count());
    best = std::min(best, std::chrono::duration<double>(t3 - t0).count());
  }
  return best;
}
int oracle_num_threads() { return omp_get_max_threads(); }
// torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU baseline sets its thread count explicitly
void oracle_set_num_threads(int32_t n) { if (n > 0) omp_set_num_threads(n); }

}  // extern "C"
