// CUDA kernels of the B200-native LM back end (sm_100a, FP64). See DESIGN.md for the data layout and the
// per-kernel roofline. Every kernel is gated by the device-resident LM state so that a whole solve can be
// enqueued without host round trips (K9, device-side controller).
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/bsnls.h"
#include "device_math.cuh"

namespace bsnls {

constexpr int kJRec = 16;   // per observation, one 128-byte line: E[6] (2x3 wrt landmark) | r[2] | Fq[6] (2x3 wrt the q tangent) | ts | 0
                            // the 2x3 block wrt t is ts * E with ts = -1 (0 when t is held constant): see reprojection_evaluate_tab
constexpr int kVRec = 18;   // per observation: V (6x3 = F^T E C), 144 bytes. c = V z is formed where it is needed from lm_z
constexpr int kGfMaxBlocks = 10, kGfMaxRes = 15, kGfMaxAmbient = 32;
constexpr int kChunkSlots = 22;        // poses per chunk: 22*23/2 = 253 pair blocks <= 256 threads
constexpr int kChunkMinObs = 100000;   // below this many observations per shard the item-list path is used instead
constexpr int kChunkMaxObs = 4096;     // observations per chunk (load balance vs. partial-block traffic)
constexpr int kChunkMaxLm = 4096;      // landmarks per chunk (bounds the stage table of chunk_syrk_kernel)
constexpr int kChunkSlack = 2;          // a chunk closes when its pose set exceeds (longest track in it) + slack: keeps most pairs present

struct LMState {
  // trust region (ceres LevenbergMarquardtStrategy)
  double radius, decrease_factor;
  double x_cost, cand_cost, model_cost_change, x_norm, cand_norm, step_norm, grad_max, grad_norm, fixed_cost;
  // pending record of an accepted step, completed once the new Jacobian / gradient exist
  double pend_cost_change, pend_step_norm, pend_relative_decrease;
  int pend_linear_iters;
  int iteration, done, need_jacobian, scale_ready, accepted, lin_fail, num_invalid;
  int termination, message_id, num_records, num_successful, num_unsuccessful, pcg_iters;
  int max_records;
};
enum MessageId { MSG_NONE = 0, MSG_MAX_ITER, MSG_GRAD_TOL, MSG_MIN_RADIUS, MSG_PARAM_TOL, MSG_FUNC_TOL,
                 MSG_INVALID_STEPS, MSG_MAX_TIME, MSG_NONFINITE };

// reduced scalars (written by reduce_kernel)
enum RedSlot { RED_COST = 0, RED_CAND_COST, RED_MODEL, RED_STEP_SQ, RED_CAND_SQ, RED_GRAD_SQ, RED_GRAD_MAX,
               RED_X_SQ, RED_FIXED, RED_NUM };

struct DeviceProblem {
  // parameter blocks
  int num_blocks, num_active; const int* blk_off; const int* blk_size; const int* blk_tan; const int* blk_manifold;
  const int* active_blocks;  // indices of active blocks
  int ne, nf, ne_dim, nf_dim, ntan;
  // camera poses = (q block, t block) pairs referenced by reprojection factors
  int num_poses; const int* pose_q_off; const int* pose_t_off; const int* pose_q_tan; const int* pose_t_tan;
  const int* pose_ptr; const int* pose_obs;   // observations of each pose (CSR), sorted observation ids
  // observations sorted by landmark (elimination order)
  int num_cams;
  int num_obs; const int* obs_pose; const int* obs_lm; const int* obs_cam; const double* obs_pixel;
  const int* lm_ptr; const int* lm_off;       // per landmark: first observation, value offset
  const CameraModel* cams;
  // generic (non-reprojection) factors, type-major
  int num_gf; const int* gf_type; const int* gf_nblk; const int* gf_nres; const int* gf_blocks;  // 10 per factor
  const long long* gf_coff; const double* gf_consts; const int* gf_loss_type; const double* gf_loss_a;
  const int* gf_joff; const int* gf_roff;     // offsets into gfJ (nres x 3*nblk row-major) and gfR
  // S pair blocks (pose a <= pose b) and their items
  int num_pairs; const int* pair_a; const int* pair_b; const int* pair_ptr; const int* item_i; const int* item_j;
  // chunks: landmark ranges accumulated on chip (chunk_syrk_kernel); per observation the slot of its pose in its chunk
  // (255: landmark handled through the item list); per pair the partial blocks the chunks produced
    // pose-major evaluation: items (pose, begin, end, 0) over the pose-sorted observation list and that list's inputs
  int num_eval_items; const int* obs_pm /* position of a (landmark-major) observation in the pose-major list = its Jacobian record */; const int4* eval_items; const int* eval_item_ptr; const int* pm_lm_off; const int* pm_cam; const double* pm_pixel;
  int chunk_max_slots; const int* chunk_pose /* kChunkSlots per chunk: pose of each slot */; const int* chunk_stage_ptr /* per chunk: first entry of stage_lm (one more entry than stages) */; const int* stage_lm /* first landmark of each stage */;
  int num_chunks; const unsigned short* obs_slot /* slot | (landmark - chunk begin) % 16 << 8; slot 255 = not in a chunk */; const int* syrk_plan; const int* chunk_begin; const int* chunk_end; const int* chunk_nslots;
  const int* pp_ptr;
  // tile format of the tensor-core chunk kernel: raw 8x8 tiles of sum W W^T per chunk; per partial-list entry the
  // chunk's tile offset and (slot A | slot B << 8 | slots << 16)
  const long long* chunk_tile_off; const long long* pp_toff; const int* pp_info;
  // generic-factor contributions: per f-block (gradient / column norms) and per 3x3 destination of S
  const int* fg_ptr; const int* fg_fac; const int* fg_lb;            // CSR over f-blocks (nf)
  int num_gdest; const int* gd_row; const int* gd_col; const int* gd_ptr; const int* gd_fac; const int* gd_bi; const int* gd_bj;
  // block-sparse reduced system (3x3 blocks, full symmetric pattern, CSR over f-blocks); sparse == 0 -> dense S
  int sparse; long long nnzb; const int* bsr_rowptr; const int* bsr_col; const int* diag_slot;
  const int* pair_slot;  // 8 per pair: slots of (qa,qb),(qa,tb),(ta,qb),(ta,tb) then of their transposes (-1: none)
  const int* gd_slot;    // 2 per destination: slot of (row,col) and of (col,row)
};

struct Options { double min_lm_diagonal, max_lm_diagonal, function_tolerance, gradient_tolerance,
                 parameter_tolerance, max_radius, min_radius, min_relative_decrease; int max_num_iterations,
                 max_invalid, jacobi_scaling; };

__device__ __forceinline__ bool gate_run(const LMState* st) { return !st->done; }
__device__ __forceinline__ bool gate_jac(const LMState* st) { return !st->done && st->need_jacobian; }

// FP64 tensor-core MMA (DMMA): D(8x8) += A(8x4, row) * B(4x8, col). Lane fragments: a = A[lane/4][lane%4],
// b = B[lane%4][lane/4], d0/d1 = D[lane/4][2*(lane%4) + {0,1}].
__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// deterministic block reduction (fixed tree), result valid in thread 0
template <int BLOCK> __device__ __forceinline__ double block_sum(double v, double* sm) {
  sm[threadIdx.x] = v; __syncthreads();
#pragma unroll
  for (int s = BLOCK / 2; s > 0; s >>= 1) { if (threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s]; __syncthreads(); }
  return sm[0];
}
template <int BLOCK> __device__ __forceinline__ double block_max(double v, double* sm) {
  sm[threadIdx.x] = v; __syncthreads();
#pragma unroll
  for (int s = BLOCK / 2; s > 0; s >>= 1) { if (threadIdx.x < s) sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + s]); __syncthreads(); }
  return sm[0];
}

// ===================================================================================================
// K4: candidate cost. Reprojection cost only at x_cand, one thread per observation of the landmark-sorted list.
// ===================================================================================================
__global__ void __launch_bounds__(256) eval_reproj_cost_kernel(DeviceProblem P, const double* __restrict__ x,
                                                               double* __restrict__ cost_part, const LMState* st) {
  if (!gate_run(st)) return;
  __shared__ double sm[256];
  const int o = blockIdx.x * 256 + threadIdx.x;
  double cost = 0.0;
  if (o < P.num_obs) {
    const int pose = P.obs_pose[o], lm = P.obs_lm[o];
    const CameraModel cam = P.cams[P.obs_cam[o]];
    const double2 px = reinterpret_cast<const double2*>(P.obs_pixel)[o];
    const double pixel[2] = {px.x, px.y};
    const int qo = P.pose_q_off[pose], to = P.pose_t_off[pose], po = P.lm_off[lm];
    const double q[4] = {x[qo], x[qo + 1], x[qo + 2], x[qo + 3]};
    const double t[3] = {x[to], x[to + 1], x[to + 2]};
    const double Pw[3] = {x[po], x[po + 1], x[po + 2]};
    cost = reprojection_cost_only(cam, pixel, q, t, Pw);
  }
  const double tot = block_sum<256>(cost, sm);
  if (threadIdx.x == 0) cost_part[blockIdx.x] = tot;
}

// ===================================================================================================
// K1 (pose-major): the Jacobian pass. Observations are walked in pose order: a warp takes up to kEvalSeg observations of
// ONE pose (host-built items), so everything that depends on the pose alone comes from the pose table (pose_prep_kernel)
// and the pose's diagonal data  H_a = sum F^T F (21 unique), g_a = sum F^T r  is reduced in the same pass with warp
// shuffles instead of re-reading the Jacobian records. Per-item partials are summed per pose by pose_reduce_kernel in item
// order: deterministic, no atomics. Jacobian records are written at their landmark-major position as before.
// ===================================================================================================
constexpr int kEvalSeg = 256, kEvalSmemCams = 8;
static_assert(kPoseTab + kEvalSmemCams * BSNLS_CAMERA_STRIDE <= 256, "one staging pass");
__global__ void __launch_bounds__(128) pose_prep_kernel(DeviceProblem P, const double* __restrict__ x, double* __restrict__ tab, const LMState* st) {
  if (!gate_jac(st)) return;
  const int a = blockIdx.x * 128 + threadIdx.x;
  if (a >= P.num_poses) return;
  const int qo = P.pose_q_off[a], to = P.pose_t_off[a];
  const double q[4] = {x[qo], x[qo + 1], x[qo + 2], x[qo + 3]};
  const double t[3] = {x[to], x[to + 1], x[to + 2]};
  double T[kPoseTab]; pose_table_build(q, t, T);
  double2* out = reinterpret_cast<double2*>(tab + (size_t)a * kPoseTab);
#pragma unroll
  for (int k = 0; k < kPoseTab / 2; ++k) out[k] = make_double2(T[2 * k], T[2 * k + 1]);
}
// One CTA per item = up to 256 consecutive observations of one pose in the pose-major list; one observation per thread.
// Jacobian records are written at the pose-major position (contiguous per CTA); consumers go through obs_pm.
// The pose's sums  sum F^T F (21 unique)  and  sum F^T r (6)  over the CTA's observations are one small Gram matrix:
// stack the rows [F | r] (2 per observation, 7 wide, padded to 8) into A (512 x 8) and form A^T A on the FP64 tensor
// cores. For mma.m8n8k4 the A^T fragment and the A fragment of a k-step are the same register (a[i][k] = b[k][i]), so a
// warp spends one shared-memory load and one DMMA per four rows; the eight warps' tiles are added in warp order.
// Deterministic (fixed order), and it frees the 32 registers per thread a shuffle reduction of 27 values would hold.
#ifndef EVAL_MINB
#define EVAL_MINB 3
#endif
__global__ void __launch_bounds__(256, EVAL_MINB) eval_reproj_pose_kernel(DeviceProblem P, const double* __restrict__ x, const double* __restrict__ tab,
                                                               double* __restrict__ Jrec, double* __restrict__ cost_part,
                                                               double* __restrict__ pose_part, const LMState* st) {
  if (!gate_jac(st)) return;
  __shared__ double sm[8][66];
  __shared__ double sCost[8];
  __shared__ double2 sOut[256 * 9];                                      // Jacobian records of the CTA, pitch 9 (conflict-free)
  __shared__ __align__(16) double sT[kPoseTab];                          // the pose's table: read by every thread (broadcast)
  __shared__ __align__(16) double sCam[kEvalSmemCams * BSNLS_CAMERA_STRIDE];   // the first few cameras (usually all of them)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int4 iv = P.eval_items[blockIdx.x];   // pose, begin, end in the pose-major list
  const int a = iv.x, it = iv.y + threadIdx.x;
  // the observation's own inputs first: the landmark point is a dependent gather (offset, then three doubles somewhere in
  // x), whose latency then overlaps the staging of the pose table and its barrier instead of following them
  const bool live = it < iv.z;
  int po = 0, ci = 0; double2 px = make_double2(0.0, 0.0);
  if (live) { po = P.pm_lm_off[it]; ci = P.pm_cam[it]; px = reinterpret_cast<const double2*>(P.pm_pixel)[it]; }
  const int qo = P.pose_q_off[a];
  const int tq = P.pose_q_tan[a], tt = P.pose_t_tan[a];
  double Pw[3] = {0.0, 0.0, 0.0};
  if (live) { Pw[0] = x[po]; Pw[1] = x[po + 1]; Pw[2] = x[po + 2]; }
  const double q[4] = {x[qo], x[qo + 1], x[qo + 2], x[qo + 3]};
  if (threadIdx.x < kPoseTab) sT[threadIdx.x] = tab[(size_t)a * kPoseTab + threadIdx.x];
  else if (threadIdx.x - kPoseTab < min(P.num_cams, kEvalSmemCams) * BSNLS_CAMERA_STRIDE)
    sCam[threadIdx.x - kPoseTab] = reinterpret_cast<const double*>(P.cams)[threadIdx.x - kPoseTab];
  __syncthreads();
  double cost = 0.0;
  if (live) {
    const double* T = sT;
    const CameraModel cam = ci < kEvalSmemCams ? reinterpret_cast<const CameraModel*>(sCam)[ci] : P.cams[ci];
    const double pixel[2] = {px.x, px.y};
    double r[2], Jq[6], JP[6];
    cost = reprojection_evaluate_tab(cam, pixel, q, T, Pw, r, Jq, JP);
    if (tq < 0) { for (int k = 0; k < 6; ++k) Jq[k] = 0.0; }
    const double ts = tt < 0 ? 0.0 : -1.0;   // d r / d t = ts * d r / d P
    // the record goes through shared memory so that the CTA's records (contiguous in the pose-major list) leave as
    // whole 512-byte warp stores instead of 32 separate lines per store instruction
    double2* out = sOut + threadIdx.x * 9;
    out[0] = make_double2(JP[0], JP[1]); out[1] = make_double2(JP[2], JP[3]); out[2] = make_double2(JP[4], JP[5]);
    out[3] = make_double2(r[0], r[1]);
    out[4] = make_double2(Jq[0], Jq[1]); out[5] = make_double2(Jq[2], Jq[3]); out[6] = make_double2(Jq[4], Jq[5]);
    out[7] = make_double2(ts, 0.0);
  } else {
    double2* out = sOut + threadIdx.x * 9;
#pragma unroll
    for (int k = 0; k < 8; ++k) out[k] = make_double2(0.0, 0.0);   // rows of zeros in the Gram matrix below
  }
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) cost += __shfl_xor_sync(0xffffffffu, cost, m);
  if (lane == 0) sCost[warp] = cost;
  __syncwarp();
  {   // Gram matrix of this warp's 64 rows [F | r | 0], read out of its records (E 0..5 | r 6,7 | Fq 8..13 | ts 14): row R of
      // the warp = residual row (R & 1) of its observation R >> 1; column g < 3: Fq, 3..5: ts E, 6: r.  d[g][2 tg], d[g][2 tg + 1]
    const int g = lane >> 2, tg = lane & 3;
    // branch-free column select: element offset in the record for residual row 0 / 1, and whether it is scaled by ts
    const int off0 = g < 3 ? 8 + g : g < 6 ? g - 3 : g == 6 ? 6 : 15, off1 = g < 3 ? 11 + g : g < 6 ? g : g == 6 ? 7 : 15;
    const bool scaled = g >= 3 && g < 6;
    double d0 = 0.0, d1 = 0.0, e0 = 0.0, e1 = 0.0;   // two accumulator pairs: halves the dependent DMMA chain
    const double* recs = reinterpret_cast<const double*>(sOut + (size_t)(32 * warp) * 9);
    const int offh = (tg & 1) ? off1 : off0;          // row R = 4 ks + tg: residual row R & 1 = tg & 1, observation 2 ks + (tg >> 1)
    const double* rec = recs + (tg >> 1) * 18;
#pragma unroll
    for (int ks = 0; ks < 16; ks += 2) {
      const double va = rec[(2 * ks) * 18 + offh] * (scaled ? rec[(2 * ks) * 18 + 14] : 1.0);
      const double vb = rec[(2 * ks + 2) * 18 + offh] * (scaled ? rec[(2 * ks + 2) * 18 + 14] : 1.0);
      dmma_m8n8k4(d0, d1, va, va);
      dmma_m8n8k4(e0, e1, vb, vb);
    }
    d0 += e0; d1 += e1;
    sm[warp][8 * g + 2 * tg] = d0; sm[warp][8 * g + 2 * tg + 1] = d1;
  }
  __syncthreads();
  {
    const int nrec = min(256, iv.z - iv.y);
    double2* dst = reinterpret_cast<double2*>(Jrec + (size_t)iv.y * kJRec);
#pragma unroll
    for (int u = 0; u < 8; ++u) { const int e = threadIdx.x + 256 * u; if ((e >> 3) < nrec) dst[e] = sOut[(e >> 3) * 9 + (e & 7)]; }
  }
  if (threadIdx.x < 28) {
    // entry k of the pose data: 0..20 = upper triangle of F^T F row-major, 21..26 = F^T r, 27 = cost
    int src = 0;
    if (threadIdx.x < 21) { int i = 0, k = threadIdx.x; while (k >= 6 - i) { k -= 6 - i; ++i; } src = 8 * i + i + k; }
    else if (threadIdx.x < 27) src = 8 * (threadIdx.x - 21) + 6;
    double tot = 0.0;
    if (threadIdx.x < 27) {
#pragma unroll
      for (int w = 0; w < 8; ++w) tot += sm[w][src];
      pose_part[(size_t)blockIdx.x * 27 + threadIdx.x] = tot;
    } else {
#pragma unroll
      for (int w = 0; w < 8; ++w) tot += sCost[w];
      cost_part[blockIdx.x] = tot;
    }
  }
}
// part_idx (optional): the pose's partials are the listed entries of pose_part (chunk-major pass: one per (chunk, slot)).
__global__ void __launch_bounds__(128) pose_reduce_kernel(DeviceProblem P, const double* __restrict__ pose_part, const int* __restrict__ part_ptr, const int* __restrict__ part_idx,
                                                          double* __restrict__ pose_H, double* __restrict__ grad, double* __restrict__ colsq, const LMState* st) {
  if (!gate_jac(st)) return;
  const int tid = blockIdx.x * 128 + threadIdx.x;
  const int a = tid / 27, k = tid - a * 27;
  if (a >= P.num_poses) return;
  double sum = 0.0;
  for (int it = part_ptr[a]; it < part_ptr[a + 1]; ++it) sum += pose_part[(size_t)(part_idx ? part_idx[it] : it) * 27 + k];
  const int tq = P.pose_q_tan[a], tt = P.pose_t_tan[a];
  if (k < 21) {
    pose_H[(size_t)a * 21 + k] = sum;
    // diagonal entries of the 21-packed upper triangle: index of (i,i) = i*6 - i(i-1)/2 = 0, 6, 11, 15, 18, 20
    const int i = k == 0 ? 0 : k == 6 ? 1 : k == 11 ? 2 : k == 15 ? 3 : k == 18 ? 4 : k == 20 ? 5 : -1;
    if (i >= 0 && i < 3 && tq >= 0) colsq[tq - P.ne_dim + i] = sum;
    if (i >= 3 && tt >= 0) colsq[tt - P.ne_dim + i - 3] = sum;
  } else {
    const int i = k - 21;
    if (i < 3 && tq >= 0) grad[tq - P.ne_dim + i] = sum;
    if (i >= 3 && tt >= 0) grad[tt - P.ne_dim + i - 3] = sum;
  }
}

// ===================================================================================================
// K2 / K3 (+K4): generic factors (IMU preintegration, 15-DoF prior, relative pose (+extrinsics), pose prior,
// ...). One warp per factor; lane k carries d/d(ambient parameter k). MODE as above.
// ===================================================================================================
template <int MODE>
__global__ void __launch_bounds__(128) eval_generic_kernel(DeviceProblem P, const double* __restrict__ x,
                                                           double* __restrict__ gfJ, double* __restrict__ gfR,
                                                           double* __restrict__ cost_part, const LMState* st) {
  if (MODE == 0 ? !gate_jac(st) : !gate_run(st)) return;
  __shared__ double sJ[4][kGfMaxRes * kGfMaxAmbient];
  __shared__ double sR[4][kGfMaxRes + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int f = blockIdx.x * 4 + warp;
  if (f >= P.num_gf) return;
  const int type = P.gf_type[f], nb = P.gf_nblk[f], nr = P.gf_nres[f];
  const int* blocks = P.gf_blocks + f * kGfMaxBlocks;
  const double* c = P.gf_consts + P.gf_coff[f];
  // gather the ambient parameter vector
  double pv[kGfMaxAmbient];
  int boff[kGfMaxBlocks + 1]; boff[0] = 0;
#pragma unroll
  for (int b = 0; b < kGfMaxBlocks; ++b) {
    if (b < nb) {
      const int blk = blocks[b], sz = P.blk_size[blk], off = P.blk_off[blk];
      for (int k = 0; k < sz; ++k) pv[boff[b] + k] = x[off + k];
      boff[b + 1] = boff[b] + sz;
    } else boff[b + 1] = boff[b];
  }
  const int namb = boff[nb];
  double* J = sJ[warp]; double* R = sR[warp];
  if (MODE == 0) {
    Dual p[kGfMaxAmbient], r[kGfMaxRes];
#pragma unroll
    for (int k = 0; k < kGfMaxAmbient; ++k) p[k] = Dual(k < namb ? pv[k] : 0.0, k == lane ? 1.0 : 0.0);
    EvalGenericFactor<Dual>(type, c, p, r);
    for (int i = 0; i < nr; ++i) { J[i * kGfMaxAmbient + lane] = (lane < namb) ? r[i].v : 0.0; if (lane == 0) R[i] = r[i].a; }
  } else {
    if (lane == 0) { double r[kGfMaxRes]; EvalGenericFactor<double>(type, c, pv, r); for (int i = 0; i < nr; ++i) R[i] = r[i]; }
  }
  __syncwarp();
  // loss
  double s = 0.0;
  for (int i = 0; i < nr; ++i) s += R[i] * R[i];
  const int lt = P.gf_loss_type[f];
  double rho[3] = {s, 1.0, 0.0};
  if (lt != 0) loss_evaluate(lt, P.gf_loss_a[f], s, rho);
  if (lane == 0) cost_part[f] = 0.5 * rho[0];
  if (MODE != 0) return;
  double sr = 1.0, rs = 1.0, asn = 0.0;
  if (lt != 0) corrector(s, rho, &sr, &rs, &asn);
  // ambient -> tangent (x plus-Jacobian for quaternion blocks), then the loss correction, then store.
  // tangent column t = 3*b + k (every block on this path has tangent size 3)
  double* Jout = gfJ + P.gf_joff[f];
  const int ncol = 3 * nb;
  if (lane < ncol) {
    const int b = lane / 3, k = lane % 3;
    const int blk = blocks[b];
    double col[kGfMaxRes];
    const bool constant = P.blk_tan[blk] < 0;
    if (P.blk_manifold[blk] == BSNLS_MANIFOLD_QUATERNION) {
      double Pj[12]; quat_plus_jacobian(pv + boff[b], Pj);
      for (int i = 0; i < nr; ++i) {
        double acc = 0.0;
        for (int a = 0; a < 4; ++a) acc += J[i * kGfMaxAmbient + boff[b] + a] * Pj[3 * a + k];
        col[i] = acc;
      }
    } else {
      for (int i = 0; i < nr; ++i) col[i] = J[i * kGfMaxAmbient + boff[b] + k];
    }
    if (constant) for (int i = 0; i < nr; ++i) col[i] = 0.0;
    if (lt != 0) {
      if (asn == 0.0) { for (int i = 0; i < nr; ++i) col[i] *= sr; }
      else {
        double rtj = 0.0;
        for (int i = 0; i < nr; ++i) rtj += col[i] * R[i];
        for (int i = 0; i < nr; ++i) col[i] = sr * (col[i] - asn * R[i] * rtj);
      }
    }
    for (int i = 0; i < nr; ++i) Jout[i * ncol + lane] = col[i];
  }
  __syncwarp();
  if (lane < nr) gfR[P.gf_roff[f] + lane] = R[lane] * rs;
}

// ===================================================================================================
// K5a: per-landmark elimination. G lanes per landmark; E^T E (3x3) and E^T r are reduced with warp shuffles.
// Outputs per landmark: Minv (6, symmetric (E^T E + D^2)^-1), g_e = E^T r, c_e = diag(E^T E), Jacobi scale;
// per observation: V = F^T E C (6x3, Minv = C C^T) and c = V (C^T g_e) so that S -= sum V_a V_b^T, rhs -= c.
// ===================================================================================================
#ifndef LM_MINB
#define LM_MINB 3
#endif
template <int G>
__global__ void __launch_bounds__(256, LM_MINB) landmark_kernel(DeviceProblem P, const double* __restrict__ Jrec,
                                                       double* __restrict__ Vrec, double* __restrict__ lm_minv,
                                                       double* __restrict__ grad, double* __restrict__ colsq,
                                                       double* __restrict__ scale, double* __restrict__ lm_z, Options opt, LMState* st) {
  if (!gate_run(st)) return;
  const int gtid = blockIdx.x * 256 + threadIdx.x;
  const int l = gtid / G, sub = gtid % G;
  const bool valid = l < P.ne;
  const int start = valid ? P.lm_ptr[l] : 0, end = valid ? P.lm_ptr[l + 1] : 0;
  double a00 = 0, a01 = 0, a02 = 0, a11 = 0, a12 = 0, a22 = 0, g0 = 0, g1 = 0, g2 = 0;
  // A lane's observations are independent gathers (index, then a 128-byte record somewhere in Jrec). A lane reading its
  // own record with eight 16-byte loads makes every load instruction of the warp touch 32 different lines (32 L1 wavefronts
  // for 512 bytes; ncu: the L1 data pipe was the busiest unit at 67 %, lg_throttle stalls). Instead the WARP moves records:
  // lane q handles 16-byte piece q % 8 of the record of lane q / 8 (+4 per step), so an instruction covers four whole
  // lines; the pieces go through a per-warp shared-memory buffer (pitch 144 bytes: conflict-free 16-byte accesses) from
  // which every lane takes its own record. The V records leave the same way: written to the buffer by their lanes,
  // stored by the warp in 16-byte pieces that are consecutive in memory (a landmark's records are contiguous in Vrec).
  // Batches of kLmBatch observations per lane, record positions kept for the second pass.
  constexpr int kLmBatch = 3;
  __shared__ double2 sRec[8][32 * 9];
  const int lane = threadIdx.x & 31;
  double2* const buf = sRec[threadIdx.x >> 5];
  int pm0[kLmBatch];
#pragma unroll
  for (int u = 0; u < kLmBatch; ++u) { const int o = start + sub + u * G; pm0[u] = o < end ? P.obs_pm[o] : -1; }
  // the first `pieces` 16-byte pieces of the 32 records at positions pm (one per lane, -1: none) -> buf[lane * 9 + piece]
  auto stage = [&](int pm, int pieces_log2) {
    const int per = 1 << pieces_log2;                       // pieces per record: 4 (E, r) or 8 (the whole record)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if ((j << 5) >= (32 << pieces_log2)) break;
      const int q = lane + 32 * j, slot = q >> pieces_log2, piece = q & (per - 1);
      const int spm = __shfl_sync(0xffffffffu, pm, slot);
      buf[slot * 9 + piece] = spm >= 0 ? reinterpret_cast<const double2*>(Jrec + (size_t)spm * kJRec)[piece] : make_double2(0.0, 0.0);
    }
    __syncwarp();
  };
#pragma unroll
  for (int u = 0; u < kLmBatch; ++u) {
    stage(pm0[u], 2);
    const double2 e0 = buf[lane * 9], e1 = buf[lane * 9 + 1], e2 = buf[lane * 9 + 2], r = buf[lane * 9 + 3];
    __syncwarp();
    // E row0 = (e0.x, e0.y, e1.x), row1 = (e1.y, e2.x, e2.y); records of missing observations are zero
    a00 += e0.x * e0.x + e1.y * e1.y; a01 += e0.x * e0.y + e1.y * e2.x; a02 += e0.x * e1.x + e1.y * e2.y;
    a11 += e0.y * e0.y + e2.x * e2.x; a12 += e0.y * e1.x + e2.x * e2.y; a22 += e1.x * e1.x + e2.y * e2.y;
    g0 += e0.x * r.x + e1.y * r.y; g1 += e0.y * r.x + e2.x * r.y; g2 += e1.x * r.x + e2.y * r.y;
  }
  for (int o = start + sub + kLmBatch * G; o < end; o += G) {   // tracks longer than kLmBatch * G
    const double2* rec = reinterpret_cast<const double2*>(Jrec + (size_t)P.obs_pm[o] * kJRec);
    const double2 e0 = rec[0], e1 = rec[1], e2 = rec[2], r = rec[3];
    a00 += e0.x * e0.x + e1.y * e1.y; a01 += e0.x * e0.y + e1.y * e2.x; a02 += e0.x * e1.x + e1.y * e2.y;
    a11 += e0.y * e0.y + e2.x * e2.x; a12 += e0.y * e1.x + e2.x * e2.y; a22 += e1.x * e1.x + e2.y * e2.y;
    g0 += e0.x * r.x + e1.y * r.y; g1 += e0.y * r.x + e2.x * r.y; g2 += e1.x * r.x + e2.y * r.y;
  }
#pragma unroll
  for (int m = 1; m < G; m <<= 1) {
    a00 += __shfl_xor_sync(0xffffffffu, a00, m); a01 += __shfl_xor_sync(0xffffffffu, a01, m);
    a02 += __shfl_xor_sync(0xffffffffu, a02, m); a11 += __shfl_xor_sync(0xffffffffu, a11, m);
    a12 += __shfl_xor_sync(0xffffffffu, a12, m); a22 += __shfl_xor_sync(0xffffffffu, a22, m);
    g0 += __shfl_xor_sync(0xffffffffu, g0, m); g1 += __shfl_xor_sync(0xffffffffu, g1, m);
    g2 += __shfl_xor_sync(0xffffffffu, g2, m);
  }
  // lanes without a landmark stay for the warp-wide record moves of the second pass (their records are empty)
  const int te = 3 * l;  // e-blocks occupy the first ne_dim tangent entries, 3 each
  double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0;
  if (valid) {
    double s0, s1, s2;
    if (!st->scale_ready) {
      s0 = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(a00)) : 1.0; s1 = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(a11)) : 1.0;
      s2 = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(a22)) : 1.0;
      if (sub == 0) { scale[te] = s0; scale[te + 1] = s1; scale[te + 2] = s2; }
    } else { s0 = scale[te]; s1 = scale[te + 1]; s2 = scale[te + 2]; }
    // LM diagonal in unscaled coordinates: D^2 = clamp(c s^2) / (radius s^2)
    const double radius = st->radius;
    const double d0 = fmin(fmax(a00 * s0 * s0, opt.min_lm_diagonal), opt.max_lm_diagonal) / (radius * s0 * s0);
    const double d1 = fmin(fmax(a11 * s1 * s1, opt.min_lm_diagonal), opt.max_lm_diagonal) / (radius * s1 * s1);
    const double d2 = fmin(fmax(a22 * s2 * s2, opt.min_lm_diagonal), opt.max_lm_diagonal) / (radius * s2 * s2);
    // Cholesky of M = E^T E + D^2 (lower L), Minv = L^-T L^-1, C = L^-T (upper), z = L^-1 g
    const double m00 = a00 + d0, m11 = a11 + d1, m22 = a22 + d2;
    const double l00 = sqrt(m00), l10 = a01 / l00, l20 = a02 / l00;
    const double t11 = m11 - l10 * l10; const double l11 = sqrt(t11), l21 = (a12 - l20 * l10) / l11;
    const double t22 = m22 - l20 * l20 - l21 * l21; const double l22 = sqrt(t22);
    if (!(m00 > 0.0) || !(t11 > 0.0) || !(t22 > 0.0)) { st->lin_fail = 1; }
    // Linv (lower): i00 = 1/l00, i10 = -l10 i00 / l11, i11 = 1/l11, i20 = -(l20 i00 + l21 i10)/l22, i21 = -l21 i11/l22, i22 = 1/l22
    i00 = 1.0 / l00; i11 = 1.0 / l11; i22 = 1.0 / l22;
    i10 = -l10 * i00 * i11; i21 = -l21 * i11 * i22; i20 = -(l20 * i00 + l21 * i10) * i22;
    const double z0 = i00 * g0, z1 = i10 * g0 + i11 * g1, z2 = i20 * g0 + i21 * g1 + i22 * g2;
    if (sub == 0) {
      double* mi = lm_minv + (size_t)l * 6;  // Minv = Linv^T Linv, stored (00,01,02,11,12,22)
      mi[0] = i00 * i00 + i10 * i10 + i20 * i20; mi[1] = i10 * i11 + i20 * i21; mi[2] = i20 * i22;
      mi[3] = i11 * i11 + i21 * i21; mi[4] = i21 * i22; mi[5] = i22 * i22;
      grad[te] = g0; grad[te + 1] = g1; grad[te + 2] = g2;
      colsq[te] = a00; colsq[te + 1] = a11; colsq[te + 2] = a22;
      lm_z[te] = z0; lm_z[te + 1] = z1; lm_z[te + 2] = z2;   // z = C^T E^T r: the Schur right-hand side is sum V z
    }
  }
  // pass 2: V = (F^T E) C with C = Linv^T (upper):  V[:,0] = W0 i00; V[:,1] = W0 i10 + W1 i11; V[:,2] = W0 i20 + W1 i21 + W2 i22
  auto make_v = [&](const double2 e0, const double2 e1, const double2 e2, const double2 f0, const double2 f1, const double2 f2, const double ts, double* v) {
    const double E0[3] = {e0.x, e0.y, e1.x}, E1[3] = {e1.y, e2.x, e2.y};
    double F0[6], F1[6];
    F0[0] = f0.x; F0[1] = f0.y; F0[2] = f1.x; F0[3] = ts * E0[0]; F0[4] = ts * E0[1]; F0[5] = ts * E0[2];
    F1[0] = f1.y; F1[1] = f2.x; F1[2] = f2.y; F1[3] = ts * E1[0]; F1[4] = ts * E1[1]; F1[5] = ts * E1[2];
#pragma unroll
    for (int a2 = 0; a2 < 6; ++a2) {
      const double w0 = F0[a2] * E0[0] + F1[a2] * E1[0], w1 = F0[a2] * E0[1] + F1[a2] * E1[1], w2 = F0[a2] * E0[2] + F1[a2] * E1[2];
      v[3 * a2] = w0 * i00; v[3 * a2 + 1] = w0 * i10 + w1 * i11; v[3 * a2 + 2] = w0 * i20 + w1 * i21 + w2 * i22;
    }
  };
#pragma unroll
  for (int u = 0; u < kLmBatch; ++u) {
    stage(pm0[u], 3);
    double v[18];
    {
      const double2* rec = buf + lane * 9;
      make_v(rec[0], rec[1], rec[2], rec[4], rec[5], rec[6], rec[7].x, v);
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 9; ++k) buf[lane * 9 + k] = make_double2(v[2 * k], v[2 * k + 1]);
    __syncwarp();
    const int o_mine = pm0[u] >= 0 ? start + sub + u * G : -1;
#pragma unroll
    for (int j = 0; j < 9; ++j) {           // 288 pieces of 16 bytes; piece q belongs to the record of lane q / 9
      const int q = lane + 32 * j, slot = q / 9, part = q - 9 * slot;
      const int o = __shfl_sync(0xffffffffu, o_mine, slot);
      if (o >= 0) reinterpret_cast<double2*>(Vrec + (size_t)o * kVRec)[part] = buf[q];
    }
    __syncwarp();
  }
  if (!valid) return;
  for (int o = start + sub + kLmBatch * G; o < end; o += G) {
    const double2* rec = reinterpret_cast<const double2*>(Jrec + (size_t)P.obs_pm[o] * kJRec);
    double v[18];
    make_v(rec[0], rec[1], rec[2], rec[4], rec[5], rec[6], rec[7].x, v);
    double2* out = reinterpret_cast<double2*>(Vrec + (size_t)o * kVRec);
#pragma unroll
    for (int k = 0; k < 9; ++k) out[k] = make_double2(v[2 * k], v[2 * k + 1]);
  }
}

// K6b: generic-factor contributions to the gradient and the column norms of every f-block.
// One thread per (f-block, component); fixed summation order.
__global__ void __launch_bounds__(128) gf_grad_kernel(DeviceProblem P, const double* __restrict__ gfJ,
                                                      const double* __restrict__ gfR, const double* __restrict__ pose_grad,
                                                      const double* __restrict__ pose_colsq, const unsigned char* __restrict__ is_pose_entry,
                                                      double* __restrict__ grad, double* __restrict__ colsq, double* __restrict__ scale,
                                                      Options opt, const LMState* st) {
  if (!gate_jac(st)) return;
  const int tid = blockIdx.x * 128 + threadIdx.x;
  const int fb = tid / 3, k = tid % 3;
  if (fb >= P.nf) return;
  double g = 0.0, c = 0.0;
  for (int it = P.fg_ptr[fb]; it < P.fg_ptr[fb + 1]; ++it) {
    const int f = P.fg_fac[it], lb = P.fg_lb[it];
    const int nr = P.gf_nres[f], ncol = 3 * P.gf_nblk[f];
    const double* J = gfJ + P.gf_joff[f]; const double* R = gfR + P.gf_roff[f];
    for (int i = 0; i < nr; ++i) { const double v = J[i * ncol + 3 * lb + k]; g += v * R[i]; c += v * v; }
  }
  const int t = P.ne_dim + 3 * fb + k;
  const bool pose = is_pose_entry[3 * fb + k];   // pose part: summed over all ranks' observations; other f-blocks have none
  const double cs = (pose ? pose_colsq[3 * fb + k] : 0.0) + c;
  grad[t] = (pose ? pose_grad[3 * fb + k] : 0.0) + g; colsq[t] = cs;
  if (!st->scale_ready) scale[t] = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(cs)) : 1.0;   // Jacobi scale, frozen at iteration 0
}

// gradient norms: |x - (x (+) -g)| per active block (TrustRegionMinimizer::EvaluateGradientAndJacobian).
__device__ __forceinline__ void quat_plus(const double* x, const double* d, double* out) {
  double qd[4]; AngleAxisToQuaternion<double>(d, qd); QuaternionProduct<double>(x, qd, out);
}
__global__ void __launch_bounds__(256) grad_norm_kernel(DeviceProblem P, const double* __restrict__ x,
                                                        const double* __restrict__ grad, double* __restrict__ part_sq,
                                                        double* __restrict__ part_max, int block0, const LMState* st) {
  if (!gate_jac(st)) return;
  __shared__ double sm[256];
  const int blk = blockIdx.x + block0;   // launched in two parts: landmark blocks before the exchange, f-blocks after it
  const int i = blk * 256 + threadIdx.x;
  double sq = 0.0, mx = 0.0;
  if (i < P.num_active && P.active_blocks[i] >= 0) {
    const int b = P.active_blocks[i], off = P.blk_off[b], t = P.blk_tan[b];
    if (P.blk_manifold[b] == BSNLS_MANIFOLD_QUATERNION) {
      const double ng[3] = {-grad[t], -grad[t + 1], -grad[t + 2]};
      double out[4]; quat_plus(x + off, ng, out);
      for (int k = 0; k < 4; ++k) { const double d = x[off + k] - out[k]; sq += d * d; mx = fmax(mx, fabs(d)); }
    } else {
      for (int k = 0; k < P.blk_size[b]; ++k) { const double d = x[off + k] - (x[off + k] + (-grad[t + k])); sq += d * d; mx = fmax(mx, fabs(d)); }
    }
  }
  const double ts = block_sum<256>(sq, sm); __syncthreads();
  const double tm = block_max<256>(mx, sm);
  if (threadIdx.x == 0) { part_sq[blk] = ts; part_max[blk] = tm; }
}

// ===================================================================================================
// K5b (tensor): the chunk accumulation as a symmetric rank-k update on the FP64 tensor cores. For a chunk with ns pose
// slots, stack the V records of landmark l into a (6 ns + 1) x 3 block column W_l (rows of absent slots zero, last
// row = z_l^T); then  sum_l W_l W_l^T  holds every pair block sum_l V_A V_B^T and, in its last column, sum_l V_A z_l
// (the Schur right-hand side). W is staged in shared memory 32 landmarks (96 columns) at a time; each warp owns a
// contiguous range of the upper-triangular 8x8 tiles and issues mma.sync.m8n8k4.f64 (DMMA) on them.
// Deterministic: fixed tile ownership, fixed k order.
// ===================================================================================================
constexpr int kSyrkStageLm = 16;                 // landmarks per stage
constexpr int kSyrkKB = 3 * kSyrkStageLm;        // 48 columns of W per stage = 12 k-steps
constexpr int kSyrkLD = 52;                      // row pitch in doubles: 52 mod 16 = 4 keeps the 64-bit fragment loads conflict-free
constexpr int kSyrkMaxTiles = 22;                // tiles per warp at 22 slots: ceil(153 / 8) + imbalance of the strip plan
constexpr int kSyrkMaxStages = 2 * (kChunkMaxLm / kSyrkStageLm) + 2;   // stages close at 16 landmarks or 256 observations
constexpr int kSyrkLanesPerRec = 9, kSyrkRecPerPass = 28, kSyrkInFlight = 8;   // 9 lanes x double2 = V (18 doubles); 28 records per pass of 252 threads
__host__ __device__ inline int syrk_smem_bytes(int nslots) { return 2 * (((6 * nslots + 1 + 7) >> 3) * 8) * kSyrkLD * 8; }

// Work plan: the upper-triangular tiles of row block I are cut into strips of up to 4 consecutive tiles (I, J0 .. J0+U-1),
// so that the A fragment is loaded once per strip and k-step; strips are dealt to the 8 warps longest-first onto the
// least loaded warp. One table for every tile count nt, built on the host (syrk_build_plan) and read from global memory.
constexpr int kSyrkMaxNt = (6 * kChunkSlots + 1 + 7) / 8, kSyrkMaxStrips = 12;
inline void syrk_build_plan(int* plan /* [(kSyrkMaxNt + 1) * 8 * kSyrkMaxStrips], entry = I | J0 << 8 | U << 16, 0 ends a list */) {
  for (int i = 0; i < (kSyrkMaxNt + 1) * 8 * kSyrkMaxStrips; ++i) plan[i] = 0;
  for (int nt = 1; nt <= kSyrkMaxNt; ++nt) {
    int strips[256], nstrips = 0;
    for (int I = 0; I < nt; ++I) {
      const int n = nt - I, k = (n + 3) / 4;
      for (int q = 0, J = I; q < k; ++q) { const int U = n / k + (q < n % k ? 1 : 0); strips[nstrips++] = I | (J << 8) | (U << 16); J += U; }
    }
    int load[8] = {0, 0, 0, 0, 0, 0, 0, 0}, cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int* pl = plan + nt * 8 * kSyrkMaxStrips;
    for (int U = 4; U >= 1; --U)
      for (int q = 0; q < nstrips; ++q) if ((strips[q] >> 16) == U) {
        int w = 0; for (int x = 1; x < 8; ++x) if (load[x] < load[w]) w = x;
        pl[w * kSyrkMaxStrips + cnt[w]++] = strips[q]; load[w] += U;
      }
    // level the remaining imbalance: peel single tiles off the busiest warp's strips
    for (int round = 0; round < 64; ++round) {
      int hi = 0, lo = 0;
      for (int x = 1; x < 8; ++x) { if (load[x] > load[hi]) hi = x; if (load[x] < load[lo]) lo = x; }
      if (load[hi] - load[lo] < 2 || cnt[lo] >= kSyrkMaxStrips) break;
      int q = -1; for (int x = 0; x < cnt[hi]; ++x) if ((pl[hi * kSyrkMaxStrips + x] >> 16) >= 2) q = x;
      if (q < 0) break;
      int& sp = pl[hi * kSyrkMaxStrips + q];
      const int I = sp & 255, J0 = (sp >> 8) & 255, U = sp >> 16;
      sp = I | (J0 << 8) | ((U - 1) << 16);
      pl[lo * kSyrkMaxStrips + cnt[lo]++] = I | ((J0 + U - 1) << 8) | (1 << 16);
      --load[hi]; ++load[lo];
    }
  }
}

template <int U, bool FULL>
__device__ __forceinline__ void syrk_strip(const double* __restrict__ pa, const double* __restrict__ pb, int ksteps, double (*accs)[2], int ab) {
  double acc[U][2];
#pragma unroll
  for (int u = 0; u < U; ++u) { acc[u][0] = accs[ab + u][0]; acc[u][1] = accs[ab + u][1]; }
  if (FULL) {
#pragma unroll
    for (int ks = 0; ks < kSyrkKB / 4; ++ks) {
      const double av = pa[4 * ks];
#pragma unroll
      for (int u = 0; u < U; ++u) dmma_m8n8k4(acc[u][0], acc[u][1], av, pb[u * 8 * kSyrkLD + 4 * ks]);
    }
  } else {
    for (int ks = 0; ks < ksteps; ++ks) {
      const double av = pa[4 * ks];
#pragma unroll
      for (int u = 0; u < U; ++u) dmma_m8n8k4(acc[u][0], acc[u][1], av, pb[u * 8 * kSyrkLD + 4 * ks]);
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) { accs[ab + u][0] = acc[u][0]; accs[ab + u][1] = acc[u][1]; }
}

__global__ void __launch_bounds__(256, 2) chunk_syrk_kernel(DeviceProblem P, const double* __restrict__ Vrec, const double* __restrict__ lm_z,
                                                            double* __restrict__ part, const LMState* st) {
  extern __shared__ __align__(16) double Wsm[];    // two stage buffers W[row * kSyrkLD + k]
  __shared__ int sPlan[8 * kSyrkMaxStrips];
  __shared__ int sO[kSyrkMaxStages], sL[kSyrkMaxStages];   // first observation / first landmark of every stage of this chunk
  if (!gate_run(st)) return;
  const int c = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, tg = lane & 3;
  const int ns = P.chunk_nslots[c], nrow = 6 * ns + 1, nt = (nrow + 7) >> 3, nrow8 = nt * 8;
  const int sp0 = P.chunk_stage_ptr[c], nst = P.chunk_stage_ptr[c + 1] - sp0 - 1;   // stages: <= 16 landmarks and <= 256 observations (host table)
  double* const buf0 = Wsm; double* const buf1 = Wsm + nrow8 * kSyrkLD;
  if (tid < 8 * kSyrkMaxStrips) sPlan[tid] = P.syrk_plan[nt * 8 * kSyrkMaxStrips + tid];
  for (int s = tid; s <= nst; s += 256) { sL[s] = P.stage_lm[sp0 + s]; sO[s] = P.lm_ptr[P.stage_lm[sp0 + s]]; }
  const int frag = g * kSyrkLD + tg;
  // accumulators of this warp's tiles; a group at a time in registers, the rest waits in (L1-resident) local memory
  double accs[kSyrkMaxTiles][2];
#pragma unroll
  for (int i = 0; i < kSyrkMaxTiles; ++i) { accs[i][0] = 0.0; accs[i][1] = 0.0; }
  // staging: 9 lanes per V record (6x3 row-major, one double2 each), 8 records in flight per lane, loaded one stage
  // ahead into registers and written to the other buffer after the tensor work of the current stage
  const int sub = tid % kSyrkLanesPerRec, oo = tid / kSyrkLanesPerRec;
  const int e0 = 2 * sub, e1 = 2 * sub + 1;   // element e of the record = V[e / 3][e % 3]
  const int off0 = (e0 / 3) * kSyrkLD + e0 % 3, off1 = (e1 / 3) * kSyrkLD + e1 % 3;
  double2 v[kSyrkInFlight]; int meta[kSyrkInFlight]; double zreg = 0.0;
  auto prefetch = [&](int s) {
    const int la = sL[s], o0 = sO[s], o1 = sO[s + 1];
#pragma unroll
    for (int u = 0; u < kSyrkInFlight; ++u) {
      const int o = o0 + oo + kSyrkRecPerPass * u;
      const bool ok = o < o1 && tid < kSyrkLanesPerRec * kSyrkRecPerPass;
      meta[u] = ok ? (int)P.obs_slot[o] : 255;
      v[u] = ok ? reinterpret_cast<const double2*>(Vrec + (size_t)o * kVRec)[sub] : make_double2(0.0, 0.0);
    }
    zreg = tid < 3 * (sL[s + 1] - la) ? lm_z[3 * (size_t)la + tid] : 0.0;
  };
  auto clear = [&](double* Wb) {
    for (int e = tid; e < nrow8 * (kSyrkKB / 2); e += 256) {
      const int r = e / (kSyrkKB / 2), k2 = e - r * (kSyrkKB / 2);
      reinterpret_cast<double2*>(Wb + r * kSyrkLD)[k2] = make_double2(0.0, 0.0);
    }
  };
  auto commit = [&](double* Wb, int s) {
    const int la = sL[s], o0 = sO[s], o1 = sO[s + 1];
#pragma unroll
    for (int u = 0; u < kSyrkInFlight; ++u) {
      if ((meta[u] & 255) == 255) continue;   // out of range, or a wide landmark (handled through the item list)
      double* base = Wb + 6 * (meta[u] & 255) * kSyrkLD + 3 * (meta[u] >> 8);
      base[off0] = v[u].x; base[off1] = v[u].y;
    }
    if (tid < kSyrkLanesPerRec * kSyrkRecPerPass)   // stages with more records than fit in flight (long tracks)
      for (int o = o0 + oo + kSyrkRecPerPass * kSyrkInFlight; o < o1; o += kSyrkRecPerPass) {
        const int m = P.obs_slot[o];
        if ((m & 255) == 255) continue;
        const double2 w = reinterpret_cast<const double2*>(Vrec + (size_t)o * kVRec)[sub];
        double* base = Wb + 6 * (m & 255) * kSyrkLD + 3 * (m >> 8);
        base[off0] = w.x; base[off1] = w.y;
      }
    if (tid < 3 * (sL[s + 1] - la)) Wb[(6 * ns) * kSyrkLD + tid] = zreg;
  };
  __syncthreads();
  prefetch(0); clear(buf0);
  __syncthreads();
  commit(buf0, 0);
  __syncthreads();
  for (int s = 0; s < nst; ++s) {
    const double* cur = (s & 1) ? buf1 : buf0; double* nxt = (s & 1) ? buf0 : buf1;
    const bool more = s + 1 < nst;
    if (more) { prefetch(s + 1); clear(nxt); }
    __syncthreads();
    const int nl = sL[s + 1] - sL[s];
    const int ksteps = (3 * nl + 3) >> 2;
    int ab = 0;
    for (int q = 0; q < kSyrkMaxStrips; ++q) {
      const int sp = sPlan[warp * kSyrkMaxStrips + q], U = sp >> 16;
      if (U == 0) break;
      const double* pa = cur + 8 * (sp & 255) * kSyrkLD + frag; const double* pb = cur + 8 * ((sp >> 8) & 255) * kSyrkLD + frag;
      if (ksteps == kSyrkKB / 4) {
        switch (U) {
          case 4: syrk_strip<4, true>(pa, pb, ksteps, accs, ab); break;
          case 3: syrk_strip<3, true>(pa, pb, ksteps, accs, ab); break;
          case 2: syrk_strip<2, true>(pa, pb, ksteps, accs, ab); break;
          default: syrk_strip<1, true>(pa, pb, ksteps, accs, ab); break;
        }
      } else {
        switch (U) {
          case 4: syrk_strip<4, false>(pa, pb, ksteps, accs, ab); break;
          case 3: syrk_strip<3, false>(pa, pb, ksteps, accs, ab); break;
          case 2: syrk_strip<2, false>(pa, pb, ksteps, accs, ab); break;
          default: syrk_strip<1, false>(pa, pb, ksteps, accs, ab); break;
        }
      }
      ab += U;
    }
    if (more) commit(nxt, s + 1);
    __syncthreads();
  }
  // raw tiles out, coalesced (a lane's two entries are adjacent); pair_reduce_kernel maps block elements to tile entries
  double* tout = part + P.chunk_tile_off[c];
  int ab = 0;
  for (int q = 0; q < kSyrkMaxStrips; ++q) {
    const int sp = sPlan[warp * kSyrkMaxStrips + q], U = sp >> 16, I = sp & 255, J0 = (sp >> 8) & 255;
    if (U == 0) break;
    const int t = I * nt - I * (I - 1) / 2 + (J0 - I);
    for (int u = 0; u < U; ++u, ++ab)
      reinterpret_cast<double2*>(tout + (size_t)(t + u) * 64)[lane] = make_double2(accs[ab][0], accs[ab][1]);
  }
}

// ===================================================================================================
// K1 + K5a + K5b fused ("chunk-major pass"): one CTA per chunk takes its observations from the parameter values to the
// chunk's partial tiles of the reduced system without a round trip through HBM in between:
//   stage by stage (<= 16 landmarks, <= 256 observations):
//     A1  one thread per observation: residual + Jacobian (bit-identical to eval_reproj_pose_kernel: same device function,
//         pose tables of the chunk's <= 22 poses in shared memory) -> record in shared memory (and, coalesced, to Jrec for
//         the back-substitution).  If the Jacobian is still valid (rejected step: only the damping changed) the records
//         are read back instead.
//     A2  E^T E (6) and E^T r (3) per landmark, summed in observation order
//     A3  3x3 Cholesky, inverse factor, z = C^T E^T r (one thread per landmark); landmark outputs to global
//     A4  V = F^T E C per observation straight into the stage's W block in shared memory; the pose's sum F^T F (21) and
//         sum F^T r (6) advance in registers of a fixed (slot, entry) thread, in landmark order
//     B   sum W W^T on the FP64 tensor cores (DMMA), as in chunk_syrk_kernel
//   end: raw tiles, per-(chunk, slot) pose sums and the chunk's cost go out.
// Removes the V records (144 B/observation written and read back), the second read of the Jacobian records and two
// kernel boundaries. Deterministic: fixed ownership, fixed orders, no atomics. Used when the shard has chunks and no
// wide landmark (track longer than the slot table, or a pose seeing a landmark twice); otherwise the three-kernel path.
// ===================================================================================================
constexpr int kFuseMaxObs = 256;
__host__ __device__ inline size_t fused_smem_bytes(int nslots) {
  const size_t nrow8 = (size_t)(((6 * nslots + 1 + 7) >> 3) * 8);
  return 8 * (nrow8 * kSyrkLD + (size_t)kFuseMaxObs * kJRec + (size_t)nslots * (kPoseTab + 4) + kEvalSmemCams * BSNLS_CAMERA_STRIDE + 2 * kSyrkStageLm * 9) +
         2 * (size_t)nslots * kSyrkStageLm + 4 * (size_t)nslots + 64;
}
__global__ void __launch_bounds__(256, 2) chunk_fused_kernel(DeviceProblem P, const double* __restrict__ x, const double* __restrict__ tab,
                                                             double* __restrict__ Jrec, double* __restrict__ lm_minv, double* __restrict__ grad,
                                                             double* __restrict__ colsq, double* __restrict__ scale, double* __restrict__ lm_z,
                                                             double* __restrict__ part, double* __restrict__ pose_part, double* __restrict__ cost_part,
                                                             Options opt, LMState* st, int with_system, int force_load) {
  extern __shared__ __align__(16) double fsm[];
  __shared__ int sPlan[8 * kSyrkMaxStrips];
  __shared__ double sRed[256];
  if (!gate_run(st)) return;
  const bool eval = st->need_jacobian && !force_load;
  if (!with_system && !eval) return;
  const int c = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, tg = lane & 3;
  const int ns = P.chunk_nslots[c], nrow = 6 * ns + 1, nt = (nrow + 7) >> 3, nrow8 = nt * 8;
  const int msl = P.chunk_max_slots;
  double* const Wb = fsm;
  double* const sJ = Wb + (size_t)(((6 * msl + 1 + 7) >> 3) * 8) * kSyrkLD;
  double* const sT = sJ + kFuseMaxObs * kJRec;
  double* const sQ = sT + msl * kPoseTab;
  double* const sCam = sQ + msl * 4;
  double* const sLmA = sCam + kEvalSmemCams * BSNLS_CAMERA_STRIDE;   // per landmark: E^T E (6), E^T r (3)
  double* const sLmB = sLmA + kSyrkStageLm * 9;                      // per landmark: inverse factor (6), z (3)
  unsigned short* const sAt = reinterpret_cast<unsigned short*>(sLmB + kSyrkStageLm * 9);   // [slot][landmark of the stage] -> observation of the stage
  unsigned char* const sFlag = reinterpret_cast<unsigned char*>(sAt + msl * kSyrkStageLm);   // per slot: bit 0 q free, bit 1 t free
  if (tid < 8 * kSyrkMaxStrips) sPlan[tid] = P.syrk_plan[nt * 8 * kSyrkMaxStrips + tid];
  if (tid < ns) {
    const int a = P.chunk_pose[c * kChunkSlots + tid];
    sFlag[tid] = (unsigned char)((P.pose_q_tan[a] >= 0 ? 1 : 0) | (P.pose_t_tan[a] >= 0 ? 2 : 0));
    if (eval) { const int qo = P.pose_q_off[a]; for (int k = 0; k < 4; ++k) sQ[4 * tid + k] = x[qo + k]; }
  }
  if (eval) {
    for (int e = tid; e < ns * kPoseTab; e += 256) { const int sl = e / kPoseTab; sT[e] = tab[(size_t)P.chunk_pose[c * kChunkSlots + sl] * kPoseTab + (e - sl * kPoseTab)]; }
    for (int e = tid; e < min(P.num_cams, kEvalSmemCams) * BSNLS_CAMERA_STRIDE; e += 256) sCam[e] = reinterpret_cast<const double*>(P.cams)[e];
  }
  const int frag = g * kSyrkLD + tg;
  double accs[kSyrkMaxTiles][2];
#pragma unroll
  for (int i = 0; i < kSyrkMaxTiles; ++i) { accs[i][0] = 0.0; accs[i][1] = 0.0; }
  // pose sums: item = (slot, entry of 27) -> (i, j) of the packed upper triangle of F^T F, or row i of F^T r
  double pacc[3] = {0.0, 0.0, 0.0}; int pi[3], pj[3];
#pragma unroll
  for (int u = 0; u < 3; ++u) {
    const int e = (tid + 256 * u) % 27;
    int i = 0, k = e; if (e < 21) { while (k >= 6 - i) { k -= 6 - i; ++i; } pi[u] = i; pj[u] = i + k; } else { pi[u] = e - 21; pj[u] = -1; }
  }
  double cost_acc = 0.0;
  const int sp0 = P.chunk_stage_ptr[c], nst = P.chunk_stage_ptr[c + 1] - sp0 - 1;
  __syncthreads();
  for (int s = 0; s < nst; ++s) {
    const int la = P.stage_lm[sp0 + s], lb = P.stage_lm[sp0 + s + 1], nlm = lb - la;
    const int o0 = P.lm_ptr[la], nobs = P.lm_ptr[lb] - o0;
    // A0: clear W and the slot table of the stage
    if (with_system)
      for (int e = tid; e < nrow8 * (kSyrkKB / 2); e += 256) { const int r = e / (kSyrkKB / 2), k2 = e - r * (kSyrkKB / 2); reinterpret_cast<double2*>(Wb + r * kSyrkLD)[k2] = make_double2(0.0, 0.0); }
    for (int e = tid; e < ns * kSyrkStageLm; e += 256) sAt[e] = 0xFFFF;
    if (!eval) { const double2* src = reinterpret_cast<const double2*>(Jrec + (size_t)o0 * kJRec); double2* dst = reinterpret_cast<double2*>(sJ); for (int e = tid; e < nobs * 8; e += 256) dst[e] = src[e]; }
    __syncthreads();
    // A1: residual + Jacobian of one observation per thread
    int slot = 0, lmslot = 0;
    if (tid < nobs) {
      const int o = o0 + tid, meta = P.obs_slot[o];
      slot = meta & 255; lmslot = meta >> 8;
      sAt[slot * kSyrkStageLm + lmslot] = (unsigned short)tid;
      if (eval) {
        const int po = P.lm_off[la + lmslot], ci = P.obs_cam[o];
        const CameraModel cam = ci < kEvalSmemCams ? reinterpret_cast<const CameraModel*>(sCam)[ci] : P.cams[ci];
        const double2 px = reinterpret_cast<const double2*>(P.obs_pixel)[o];
        const double pixel[2] = {px.x, px.y};
        const double Pw[3] = {x[po], x[po + 1], x[po + 2]};
        double r[2], Jq[6], JP[6];
        cost_acc += reprojection_evaluate_tab(cam, pixel, sQ + 4 * slot, sT + kPoseTab * slot, Pw, r, Jq, JP);
        const unsigned char fl = sFlag[slot];
        if (!(fl & 1)) { for (int k = 0; k < 6; ++k) Jq[k] = 0.0; }
        double2* out = reinterpret_cast<double2*>(sJ + tid * kJRec);
        out[0] = make_double2(JP[0], JP[1]); out[1] = make_double2(JP[2], JP[3]); out[2] = make_double2(JP[4], JP[5]);
        out[3] = make_double2(r[0], r[1]);
        out[4] = make_double2(Jq[0], Jq[1]); out[5] = make_double2(Jq[2], Jq[3]); out[6] = make_double2(Jq[4], Jq[5]);
        out[7] = make_double2((fl & 2) ? -1.0 : 0.0, 0.0);
      }
    }
    __syncthreads();
    if (eval) { const double2* src = reinterpret_cast<const double2*>(sJ); double2* dst = reinterpret_cast<double2*>(Jrec + (size_t)o0 * kJRec); for (int e = tid; e < nobs * 8; e += 256) dst[e] = src[e]; }
    // A2: normal equations of the stage's landmarks
    if (tid < kSyrkStageLm * 9) {
      const int lm = tid / 9, e = tid - 9 * lm;
      if (lm < nlm) {
        const int k0 = P.lm_ptr[la + lm] - o0, k1 = P.lm_ptr[la + lm + 1] - o0;
        double sum = 0.0;
        if (e < 6) {
          const int i = e < 3 ? 0 : e < 5 ? 1 : 2, j = e < 3 ? e : e < 5 ? e - 2 : 2;
          for (int k = k0; k < k1; ++k) { const double* rec = sJ + k * kJRec; sum += rec[i] * rec[j] + rec[3 + i] * rec[3 + j]; }
        } else {
          const int i = e - 6;
          for (int k = k0; k < k1; ++k) { const double* rec = sJ + k * kJRec; sum += rec[i] * rec[6] + rec[3 + i] * rec[7]; }
        }
        sLmA[tid] = sum;
      }
    }
    __syncthreads();
    // A3: one thread per landmark: damping, Cholesky, inverse factor, z; landmark outputs
    if (tid < nlm) {
      const int l = la + tid, te = 3 * l;
      const double* A = sLmA + 9 * tid;
      const double a00 = A[0], a01 = A[1], a02 = A[2], a11 = A[3], a12 = A[4], a22 = A[5], g0 = A[6], g1 = A[7], g2 = A[8];
      double s0, s1, s2;
      if (!st->scale_ready) {
        s0 = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(a00)) : 1.0; s1 = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(a11)) : 1.0;
        s2 = opt.jacobi_scaling ? 1.0 / (1.0 + sqrt(a22)) : 1.0;
        scale[te] = s0; scale[te + 1] = s1; scale[te + 2] = s2;
      } else { s0 = scale[te]; s1 = scale[te + 1]; s2 = scale[te + 2]; }
      const double radius = st->radius;
      const double d0 = fmin(fmax(a00 * s0 * s0, opt.min_lm_diagonal), opt.max_lm_diagonal) / (radius * s0 * s0);
      const double d1 = fmin(fmax(a11 * s1 * s1, opt.min_lm_diagonal), opt.max_lm_diagonal) / (radius * s1 * s1);
      const double d2 = fmin(fmax(a22 * s2 * s2, opt.min_lm_diagonal), opt.max_lm_diagonal) / (radius * s2 * s2);
      const double m00 = a00 + d0, m11 = a11 + d1, m22 = a22 + d2;
      const double l00 = sqrt(m00), l10 = a01 / l00, l20 = a02 / l00;
      const double t11 = m11 - l10 * l10; const double l11 = sqrt(t11), l21 = (a12 - l20 * l10) / l11;
      const double t22 = m22 - l20 * l20 - l21 * l21; const double l22 = sqrt(t22);
      if (!(m00 > 0.0) || !(t11 > 0.0) || !(t22 > 0.0)) { st->lin_fail = 1; }
      const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
      const double i10 = -l10 * i00 * i11, i21 = -l21 * i11 * i22, i20 = -(l20 * i00 + l21 * i10) * i22;
      const double z0 = i00 * g0, z1 = i10 * g0 + i11 * g1, z2 = i20 * g0 + i21 * g1 + i22 * g2;
      double* B = sLmB + 9 * tid;
      B[0] = i00; B[1] = i10; B[2] = i11; B[3] = i20; B[4] = i21; B[5] = i22; B[6] = z0; B[7] = z1; B[8] = z2;
      double* mi = lm_minv + (size_t)l * 6;
      mi[0] = i00 * i00 + i10 * i10 + i20 * i20; mi[1] = i10 * i11 + i20 * i21; mi[2] = i20 * i22;
      mi[3] = i11 * i11 + i21 * i21; mi[4] = i21 * i22; mi[5] = i22 * i22;
      grad[te] = g0; grad[te + 1] = g1; grad[te + 2] = g2;
      colsq[te] = a00; colsq[te + 1] = a11; colsq[te + 2] = a22;
      lm_z[te] = z0; lm_z[te + 1] = z1; lm_z[te + 2] = z2;
    }
    __syncthreads();
    // A4: V = (F^T E) C into W; z row; pose sums
    if (with_system) {
      if (tid < nobs) {
        const double* rec = sJ + tid * kJRec; const double* B = sLmB + 9 * lmslot;
        const double E0[3] = {rec[0], rec[1], rec[2]}, E1[3] = {rec[3], rec[4], rec[5]}, ts = rec[14];
        const double F0[6] = {rec[8], rec[9], rec[10], ts * E0[0], ts * E0[1], ts * E0[2]}, F1[6] = {rec[11], rec[12], rec[13], ts * E1[0], ts * E1[1], ts * E1[2]};
        const double i00 = B[0], i10 = B[1], i11 = B[2], i20 = B[3], i21 = B[4], i22 = B[5];
        double* wrow = Wb + (size_t)(6 * slot) * kSyrkLD + 3 * lmslot;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
          const double w0 = F0[a] * E0[0] + F1[a] * E1[0], w1 = F0[a] * E0[1] + F1[a] * E1[1], w2 = F0[a] * E0[2] + F1[a] * E1[2];
          wrow[a * kSyrkLD] = w0 * i00; wrow[a * kSyrkLD + 1] = w0 * i10 + w1 * i11; wrow[a * kSyrkLD + 2] = w0 * i20 + w1 * i21 + w2 * i22;
        }
      }
      if (tid < 3 * nlm) Wb[(size_t)(6 * ns) * kSyrkLD + tid] = sLmB[9 * (tid / 3) + 6 + tid % 3];
    }
    if (eval) {
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int it = tid + 256 * u;
        if (it >= 27 * ns) break;
        const unsigned short* at = sAt + (it / 27) * kSyrkStageLm;
        const int i = pi[u], j = pj[u];
        double acc = pacc[u];
        for (int m = 0; m < nlm; ++m) {
          const int t = at[m];
          if (t == 0xFFFF) continue;
          const double* rec = sJ + t * kJRec;
          // F row 0 = (Fq[0..2], ts E0), row 1 = (Fq[3..5], ts E1)
          const double ts = rec[14];
          const double f0i = i < 3 ? rec[8 + i] : ts * rec[i - 3], f1i = i < 3 ? rec[11 + i] : ts * rec[i];
          if (j >= 0) { const double f0j = j < 3 ? rec[8 + j] : ts * rec[j - 3], f1j = j < 3 ? rec[11 + j] : ts * rec[j]; acc += f0i * f0j + f1i * f1j; }
          else acc += f0i * rec[6] + f1i * rec[7];
        }
        pacc[u] = acc;
      }
    }
    __syncthreads();
    // B: sum W W^T of the stage on the tensor cores
    if (with_system) {
      const int ksteps = (3 * nlm + 3) >> 2;
      int ab = 0;
      for (int q = 0; q < kSyrkMaxStrips; ++q) {
        const int sp = sPlan[warp * kSyrkMaxStrips + q], U = sp >> 16;
        if (U == 0) break;
        const double* pa = Wb + 8 * (sp & 255) * kSyrkLD + frag; const double* pb = Wb + 8 * ((sp >> 8) & 255) * kSyrkLD + frag;
        if (ksteps == kSyrkKB / 4) {
          switch (U) {
            case 4: syrk_strip<4, true>(pa, pb, ksteps, accs, ab); break;
            case 3: syrk_strip<3, true>(pa, pb, ksteps, accs, ab); break;
            case 2: syrk_strip<2, true>(pa, pb, ksteps, accs, ab); break;
            default: syrk_strip<1, true>(pa, pb, ksteps, accs, ab); break;
          }
        } else {
          switch (U) {
            case 4: syrk_strip<4, false>(pa, pb, ksteps, accs, ab); break;
            case 3: syrk_strip<3, false>(pa, pb, ksteps, accs, ab); break;
            case 2: syrk_strip<2, false>(pa, pb, ksteps, accs, ab); break;
            default: syrk_strip<1, false>(pa, pb, ksteps, accs, ab); break;
          }
        }
        ab += U;
      }
    }
    __syncthreads();
  }
  if (with_system) {
    double* tout = part + P.chunk_tile_off[c];
    int ab = 0;
    for (int q = 0; q < kSyrkMaxStrips; ++q) {
      const int sp = sPlan[warp * kSyrkMaxStrips + q], U = sp >> 16, I = sp & 255, J0 = (sp >> 8) & 255;
      if (U == 0) break;
      const int t = I * nt - I * (I - 1) / 2 + (J0 - I);
      for (int u = 0; u < U; ++u, ++ab)
        reinterpret_cast<double2*>(tout + (size_t)(t + u) * 64)[lane] = make_double2(accs[ab][0], accs[ab][1]);
    }
  }
  if (eval) {
#pragma unroll
    for (int u = 0; u < 3; ++u) { const int it = tid + 256 * u; if (it < 27 * ns) pose_part[((size_t)c * kChunkSlots + it / 27) * 27 + it % 27] = pacc[u]; }
    const double tot = block_sum<256>(cost_acc, sRed);
    if (tid == 0) cost_part[c] = tot;
  }
}

// ===================================================================================================
// K5b: reduced-system assembly, atomics-free. One warp per pose-pair block (a <= b): lanes stride over the
// items (observation pairs sharing a landmark), each lane accumulates a full 6x6, butterfly reduction gives a
// fixed summation order. Diagonal blocks add sum F^T F and produce the reduced right-hand side.
// S is dense column-major (ld = nf_dim), lower triangle.
// ===================================================================================================
__device__ __forceinline__ void store_sub3(double* S, int ld, int rt, int ct, const double* B /*3x3 row-major (rt rows, ct cols)*/) {
  // write the 3x3 block at rows rt.., cols ct.. into the lower triangle (transposed if it lies above)
  if (rt >= ct) { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) S[(size_t)(ct + j) * ld + rt + i] = B[3 * i + j]; }
  else { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) S[(size_t)(rt + i) * ld + ct + j] = B[3 * i + j]; }
}
__global__ void __launch_bounds__(128) pair_kernel(DeviceProblem P, const double* __restrict__ Vrec,
                                                   const double* __restrict__ lm_z,
                                                   const double* __restrict__ pose_H, const double* __restrict__ grad,
                                                   double* __restrict__ S, double* __restrict__ rhs, const int* __restrict__ pair_list, int num_list, const LMState* st) {
  if (!gate_run(st)) return;
  const int idx = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (idx >= num_list) return;
  const int pb = pair_list ? pair_list[idx] : idx;
  const int a = P.pair_a[pb], b = P.pair_b[pb];
  double acc[36], cacc[6];
#pragma unroll
  for (int k = 0; k < 36; ++k) acc[k] = 0.0;
#pragma unroll
  for (int k = 0; k < 6; ++k) cacc[k] = 0.0;
  for (int it = P.pair_ptr[pb] + lane; it < P.pair_ptr[pb + 1]; it += 32) {
    const int oi = P.item_i[it], oj = P.item_j[it];
    const double2* vi = reinterpret_cast<const double2*>(Vrec + (size_t)oi * kVRec);
    const double2* vj = reinterpret_cast<const double2*>(Vrec + (size_t)oj * kVRec);
    double Vi[18], Vj[18];
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 u = vi[k]; Vi[2 * k] = u.x; Vi[2 * k + 1] = u.y; }
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 u = vj[k]; Vj[2 * k] = u.x; Vj[2 * k + 1] = u.y; }
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j)
        acc[6 * i + j] += Vi[3 * i] * Vj[3 * j] + Vi[3 * i + 1] * Vj[3 * j + 1] + Vi[3 * i + 2] * Vj[3 * j + 2];
    if (oi == oj) {   // c = V z of this observation: the Schur right-hand side
      const double* z = lm_z + 3 * (size_t)P.obs_lm[oi];
#pragma unroll
      for (int k = 0; k < 6; ++k) cacc[k] += Vi[3 * k] * z[0] + Vi[3 * k + 1] * z[1] + Vi[3 * k + 2] * z[2];
    }
  }
#pragma unroll
  for (int m = 1; m < 32; m <<= 1) {
#pragma unroll
    for (int k = 0; k < 36; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], m);
#pragma unroll
    for (int k = 0; k < 6; ++k) cacc[k] += __shfl_xor_sync(0xffffffffu, cacc[k], m);
  }
  if (lane != 0) return;
  const int ld = P.nf_dim;
  const int ta[2] = {P.pose_q_tan[a] < 0 ? -1 : P.pose_q_tan[a] - P.ne_dim, P.pose_t_tan[a] < 0 ? -1 : P.pose_t_tan[a] - P.ne_dim};
  const int tb[2] = {P.pose_q_tan[b] < 0 ? -1 : P.pose_q_tan[b] - P.ne_dim, P.pose_t_tan[b] < 0 ? -1 : P.pose_t_tan[b] - P.ne_dim};
  double blk[36];
  if (a == b) {
    const double* H = pose_H + (size_t)a * 21;
    int k = 0;
    for (int i = 0; i < 6; ++i) for (int j = i; j < 6; ++j) { const double h = H[k++]; blk[6 * i + j] = h - acc[6 * i + j]; if (j != i) blk[6 * j + i] = h - acc[6 * j + i]; }
    for (int s = 0; s < 2; ++s) if (ta[s] >= 0) for (int k2 = 0; k2 < 3; ++k2) rhs[ta[s] + k2] = -cacc[3 * s + k2];   // + gradient in finish_system
  } else {
    for (int k = 0; k < 36; ++k) blk[k] = -acc[k];
  }
  for (int sa = 0; sa < 2; ++sa)
    for (int sb = 0; sb < 2; ++sb) {
      if (ta[sa] < 0 || tb[sb] < 0) continue;
      if (a == b && sa > sb) continue;             // (t,q) duplicates (q,t)^T
      double B[9];
      for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) B[3 * i + j] = blk[6 * (3 * sa + i) + 3 * sb + j];
      if (!P.sparse) { store_sub3(S, ld, ta[sa], tb[sb], B); continue; }
      const int slot = P.pair_slot[pb * 8 + 2 * sa + sb], slot_t = P.pair_slot[pb * 8 + 4 + 2 * sa + sb];
      for (int k = 0; k < 9; ++k) S[(size_t)slot * 9 + k] = B[k];
      if (slot_t >= 0 && slot_t != slot) for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) S[(size_t)slot_t * 9 + 3 * j + i] = B[3 * i + j];
    }
}

// K5b (reduce): final reduction of the reduced-system blocks. One thread per (pose pair, block element): sums the
// chunk partials of its element in list order (coalesced: the 36 elements of a partial block are contiguous), adds the
// contributions of wide landmarks from the item list, subtracts from sum F^T F on the diagonal blocks and writes
// S (dense lower triangle or 3x3 block-sparse) and the reduced right-hand side. 4 pairs per 144-thread CTA.
// pair_list (optional): the pairs this rank's landmarks contribute to (sharded solve); the others keep their zeros.
__global__ void __launch_bounds__(144) pair_reduce_kernel(DeviceProblem P, const double* __restrict__ Vrec,
                                                          const double* __restrict__ part, const double* __restrict__ lm_z,
                                                          const double* __restrict__ pose_H, double* __restrict__ S,
                                                          double* __restrict__ rhs, const int* __restrict__ pair_list, int num_list, const LMState* st) {
  if (!gate_run(st)) return;
  const int idx = blockIdx.x * 4 + threadIdx.x / 36, e = threadIdx.x % 36;
  if (idx >= num_list) return;
  const int pb = pair_list ? pair_list[idx] : idx;
  const int a = P.pair_a[pb], b = P.pair_b[pb];
  const int i = e / 6, j = e - 6 * i;
  double acc = 0.0, cacc = 0.0;
  {
    for (int it = P.pp_ptr[pb]; it < P.pp_ptr[pb + 1]; ++it) {
      const int info = P.pp_info[it], A = info & 255, B = (info >> 8) & 255, ns = info >> 16;
      const int nt = (6 * ns + 8) >> 3;
      const double* tp = part + P.pp_toff[it];
      int r = 6 * A + i, c = 6 * B + j;
      if (r > c) { const int t = r; r = c; c = t; }          // only on diagonal blocks (A == B): symmetric
      int I = r >> 3, J = c >> 3;
      acc += tp[(size_t)(I * nt - I * (I - 1) / 2 + (J - I)) * 64 + (r & 7) * 8 + (c & 7)];
      if (e < 6 && A == B) {                                  // last column of W W^T: sum V_A z
        r = 6 * A + e; c = 6 * ns; I = r >> 3; J = c >> 3;
        cacc += tp[(size_t)(I * nt - I * (I - 1) / 2 + (J - I)) * 64 + (r & 7) * 8 + (c & 7)];
      }
    }
  }
  for (int it = P.pair_ptr[pb]; it < P.pair_ptr[pb + 1]; ++it) {   // wide landmarks
    const int oi = P.item_i[it], oj = P.item_j[it];
    const double* vi = Vrec + (size_t)oi * kVRec + 3 * i; const double* vj = Vrec + (size_t)oj * kVRec + 3 * j;
    acc += vi[0] * vj[0] + vi[1] * vj[1] + vi[2] * vj[2];
    if (e < 6 && oi == oj) { const double* z = lm_z + 3 * (size_t)P.obs_lm[oi]; const double* ve = Vrec + (size_t)oi * kVRec + 3 * e; cacc += ve[0] * z[0] + ve[1] * z[1] + ve[2] * z[2]; }
  }
  const int sa = i / 3, sb = j / 3, ii = i - 3 * sa, jj = j - 3 * sb;
  const int ta = (sa ? P.pose_t_tan[a] : P.pose_q_tan[a]), tb = (sb ? P.pose_t_tan[b] : P.pose_q_tan[b]);
  double v = -acc;
  if (a == b) {
    const int lo = min(i, j), hi = max(i, j);
    v += pose_H[(size_t)a * 21 + lo * 6 - lo * (lo - 1) / 2 + (hi - lo)];   // packed upper triangle, row-major
    if (e < 6) { const int t = (e < 3 ? P.pose_q_tan[a] : P.pose_t_tan[a]); if (t >= 0) rhs[t - P.ne_dim + e % 3] = -cacc; }   // + gradient in finish_system
  }
  if (ta < 0 || tb < 0) return;
  if (P.sparse) {
    if (a == b && sa > sb) return;   // the (q,t) threads write both the block and its transpose
    const int slot = P.pair_slot[pb * 8 + 2 * sa + sb], slot_t = P.pair_slot[pb * 8 + 4 + 2 * sa + sb];
    S[(size_t)slot * 9 + 3 * ii + jj] = v;
    if (slot_t >= 0 && slot_t != slot) S[(size_t)slot_t * 9 + 3 * jj + ii] = v;
    return;
  }
  const int r = ta - P.ne_dim + ii, c = tb - P.ne_dim + jj, ld = P.nf_dim;
  if (a == b) { if (r >= c) S[(size_t)c * ld + r] = v; }     // symmetric block: the mirrored thread holds the other triangle
  else if (r >= c) S[(size_t)c * ld + r] = v; else S[(size_t)r * ld + c] = v;
}

// Sharded solve: everything a pass exchanges travels in ONE all-reduce (sum) of a packed buffer:
//   [ lower-triangular blocks of the (symmetric) block-sparse S | rhs | pose gradient | pose column norms |
//     rank-local scalar sums | one slot per rank for the rank-local max | number of ranks whose elimination failed ]
// pack gathers it (the segments before `first` are left out in a Jacobian-only pass), unpack scatters the sums: S blocks
// to the block and its transpose, the max as the maximum over the per-rank slots.
struct ExchangeLayout { long long nS; int nf_dim, world, rank, num_sums, max_slot; const int* low_slot; const int* low_slot_t; };
__global__ void __launch_bounds__(256) pack_exchange_kernel(ExchangeLayout X, long long first, const double* __restrict__ S, const double* __restrict__ rhs,
                                                            const double* __restrict__ gradL, const double* __restrict__ colsqL, const double* __restrict__ red,
                                                            const LMState* st, double* __restrict__ out) {
  const long long e = first + (long long)blockIdx.x * 256 + threadIdx.x;
  const long long o1 = X.nS + X.nf_dim, o2 = o1 + X.nf_dim, o3 = o2 + X.nf_dim, o4 = o3 + X.num_sums, o5 = o4 + X.world;
  if (e > o5) return;
  if (e == o5) { out[e] = st->lin_fail ? 1.0 : 0.0; return; }   // a failed landmark elimination on one rank fails the step on all of them
  double v;
  if (e < X.nS) v = S[(size_t)X.low_slot[e / 9] * 9 + e % 9];
  else if (e < o1) v = rhs[e - X.nS];
  else if (e < o2) v = gradL[e - o1];
  else if (e < o3) v = colsqL[e - o2];
  else if (e < o4) v = red[e - o3];
  else v = (e - o4 == X.rank) ? red[X.max_slot] : 0.0;
  out[e] = v;
}
__global__ void __launch_bounds__(256) unpack_exchange_kernel(ExchangeLayout X, long long first, const double* __restrict__ in, double* __restrict__ S,
                                                              double* __restrict__ rhs, double* __restrict__ gradS, double* __restrict__ colsqS,
                                                              double* __restrict__ red, LMState* st) {
  const long long e = first + (long long)blockIdx.x * 256 + threadIdx.x;
  const long long o1 = X.nS + X.nf_dim, o2 = o1 + X.nf_dim, o3 = o2 + X.nf_dim, o4 = o3 + X.num_sums, o5 = o4 + X.world;
  if (e > o5) return;
  if (e == o5) { if (in[e] > 0.0) st->lin_fail = 1; return; }
  const double v = in[e];
  if (e < X.nS) {
    const int i = (int)(e / 9), w = (int)(e % 9);
    S[(size_t)X.low_slot[i] * 9 + w] = v;
    const int t = X.low_slot_t[i];
    if (t >= 0) S[(size_t)t * 9 + 3 * (w % 3) + w / 3] = v;
  } else if (e < o1) rhs[e - X.nS] = v;
  else if (e < o2) gradS[e - o1] = v;
  else if (e < o3) colsqS[e - o2] = v;
  else if (e < o4) red[e - o3] = v;
  else if (e == o4) { double m = v; for (int r = 1; r < X.world; ++r) m = fmax(m, in[o4 + r]); red[X.max_slot] = m; }
}

// K6c: generic-factor contributions to S (3x3 destinations, row >= col) and to rhs. 9 threads per destination.
__global__ void __launch_bounds__(144) gf_schur_kernel(DeviceProblem P, const double* __restrict__ gfJ,
                                                       double* __restrict__ S, const LMState* st) {
  if (!gate_run(st)) return;
  const int d = blockIdx.x * 16 + threadIdx.x / 9, e = threadIdx.x % 9;
  if (d >= P.num_gdest) return;
  const int i = e / 3, j = e % 3;
  double acc = 0.0;
  for (int it = P.gd_ptr[d]; it < P.gd_ptr[d + 1]; ++it) {
    const int f = P.gd_fac[it], bi = P.gd_bi[it], bj = P.gd_bj[it];
    const int nr = P.gf_nres[f], ncol = 3 * P.gf_nblk[f];
    const double* J = gfJ + P.gf_joff[f];
    for (int r = 0; r < nr; ++r) acc += J[r * ncol + 3 * bi + i] * J[r * ncol + 3 * bj + j];
  }
  if (P.sparse) {
    const int slot = P.gd_slot[2 * d], slot_t = P.gd_slot[2 * d + 1];
    S[(size_t)slot * 9 + 3 * i + j] += acc;
    if (slot_t != slot) S[(size_t)slot_t * 9 + 3 * j + i] += acc;
    return;
  }
  const int rt = 3 * P.gd_row[d] + i, ct = 3 * P.gd_col[d] + j;
  if (rt >= ct) S[(size_t)ct * P.nf_dim + rt] += acc;
}
// rhs of f-blocks that are not camera poses, then D_f^2 on the diagonal.
__global__ void __launch_bounds__(256) finish_system_kernel(DeviceProblem P, const double* __restrict__ grad,
                                                            const double* __restrict__ colsq, const double* __restrict__ scale,
                                                            const unsigned char* __restrict__ is_pose_entry, double* __restrict__ S,
                                                            double* __restrict__ rhs, Options opt, const LMState* st) {
  if (!gate_run(st)) return;
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= P.nf_dim) return;
  const double s = scale[P.ne_dim + j];
  const double d = fmin(fmax(colsq[P.ne_dim + j] * s * s, opt.min_lm_diagonal), opt.max_lm_diagonal) / (st->radius * s * s);
  if (P.sparse) S[(size_t)P.diag_slot[j / 3] * 9 + 4 * (j % 3)] += d; else S[(size_t)j * P.nf_dim + j] += d;
  rhs[j] = (is_pose_entry[j] ? rhs[j] : 0.0) + grad[P.ne_dim + j];
}


// ===================================================================================================
// K8: dense FP64 Cholesky of the reduced system for sliding windows (column-major, lower), one CTA. Larger systems go
// to the tile-sparse Cholesky (tile_chol.cuh).
// ===================================================================================================
// n <= 160 (a 10-keyframe VIO window has 150 unknowns): the lower triangle lives in
// registers as 4x4 blocks, two per thread, plus one extra block row that carries the right-hand side, so the forward
// substitution happens inside the factorisation sweep. Four pivots per step and two barriers per step: the owner of
// the diagonal block factors it and publishes the inverse of its 4x4 factor, the block column is multiplied by it and
// published, every block to the right gets a rank-4 update. The backward substitution walks the block columns once
// more (partial products of a column are combined by one warp). ~80 barriers instead of ~1000.
constexpr int kRbMaxN = 160, kRbNB = kRbMaxN / 4, kRbThreads = 512, kRbMaxBlocks = kRbNB * (kRbNB + 3) / 2;
struct RbBlock { double a[4][4]; int bi, bj; };   // bj < 0: unused slot
__device__ __forceinline__ void rb_panel(RbBlock& B, int jb, int nb, const double2* __restrict__ sLi, double2 (*sPan)[kRbNB + 4], double* ys) {
  if (B.bj != jb || B.bi <= jb) return;
  double Li[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { const double2 u0 = sLi[2 * i], u1 = sLi[2 * i + 1]; Li[i][0] = u0.x; Li[i][1] = u0.y; Li[i][2] = u1.x; Li[i][3] = u1.y; }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double l[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) { double acc = 0.0;
#pragma unroll
      for (int m = 0; m <= c; ++m) acc += B.a[i][m] * Li[c][m];
      l[c] = acc; }
#pragma unroll
    for (int c = 0; c < 4; ++c) B.a[i][c] = l[c];
    sPan[2 * i][B.bi] = make_double2(l[0], l[1]); sPan[2 * i + 1][B.bi] = make_double2(l[2], l[3]);
  }
  if (B.bi == nb) {   // the right-hand-side row: y of this block column is final
#pragma unroll
    for (int c = 0; c < 4; ++c) ys[4 * jb + c] = B.a[0][c];
  }
}
__device__ __forceinline__ void rb_update(RbBlock& B, int jb, const double2 (*sPan)[kRbNB + 4]) {
  if (B.bj <= jb) return;
  double Pr[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { const double2 u0 = sPan[2 * i][B.bi], u1 = sPan[2 * i + 1][B.bi]; Pr[i][0] = u0.x; Pr[i][1] = u0.y; Pr[i][2] = u1.x; Pr[i][3] = u1.y; }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const double2 u0 = sPan[2 * c][B.bj], u1 = sPan[2 * c + 1][B.bj];
#pragma unroll
    for (int i = 0; i < 4; ++i) B.a[i][c] -= Pr[i][0] * u0.x + Pr[i][1] * u0.y + Pr[i][2] * u1.x + Pr[i][3] * u1.y;
  }
}
__device__ __forceinline__ void rb_back(const RbBlock& B, int jb, int nb, const double* xs, double (*part)[4]) {
  if (B.bj != jb || B.bi <= jb || B.bi >= nb) return;
#pragma unroll
  for (int c = 0; c < 4; ++c) part[B.bi][c] = B.a[0][c] * xs[4 * B.bi] + B.a[1][c] * xs[4 * B.bi + 1] + B.a[2][c] * xs[4 * B.bi + 2] + B.a[3][c] * xs[4 * B.bi + 3];
}
__global__ void __launch_bounds__(kRbThreads, 1) chol_reg_kernel(const double* __restrict__ S, int n, const double* __restrict__ rhs,
                                                                 double* __restrict__ y, LMState* st) {
  __shared__ double2 sPan[8][kRbNB + 4];       // the block column, element-major: sPan[e][bi] (rows of a warp's blocks are adjacent)
  __shared__ double2 sLiAll[kRbNB][8];         // inverse of every 4x4 diagonal factor, row-major
  __shared__ double xs[kRbMaxN + 4], ys[kRbMaxN + 4];
  __shared__ double part[kRbNB + 1][4];
  __shared__ unsigned short sBlk[kRbMaxBlocks];
  __shared__ int failed;
  if (!gate_run(st) || st->lin_fail) return;
  const int tid = threadIdx.x, nb = (n + 3) >> 2, total = nb * (nb + 3) / 2;
  if (tid == 0) failed = 0;
  // blocks in column-major order: column bj holds rows bj..nb (row nb = the right-hand side)
  if (tid < nb) { int o = tid * (nb + 1) - tid * (tid - 1) / 2; for (int bi = tid; bi <= nb; ++bi) sBlk[o++] = (unsigned short)(bi | (tid << 8)); }
  __syncthreads();
  RbBlock B0, B1;
  auto load = [&](RbBlock& B, int L) {
    if (L >= total) { B.bi = -1; B.bj = -1; return; }
    B.bi = sBlk[L] & 255; B.bj = sBlk[L] >> 8;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = 4 * B.bi + i, cc = 4 * B.bj + c;
        double v;
        if (B.bi == nb) v = (i == 0 && cc < n) ? rhs[cc] : 0.0;
        else if (r < n && cc < n) v = r >= cc ? S[(size_t)cc * n + r] : 0.0;
        else v = r == cc ? 1.0 : 0.0;     // padding up to a multiple of four
        B.a[i][c] = v;
      }
  };
  load(B0, tid); load(B1, tid + kRbThreads);
  bool bad = false;
  auto diag = [&](RbBlock& B, int jb) {
    if (B.bi != jb || B.bj != jb) return;
    const double r0 = rsqrt(B.a[0][0]);
    const double l10 = B.a[1][0] * r0, l20 = B.a[2][0] * r0, l30 = B.a[3][0] * r0;
    const double d1 = B.a[1][1] - l10 * l10, r1 = rsqrt(d1);
    const double l21 = (B.a[2][1] - l20 * l10) * r1, l31 = (B.a[3][1] - l30 * l10) * r1;
    const double d2 = B.a[2][2] - l20 * l20 - l21 * l21, r2 = rsqrt(d2);
    const double l32 = (B.a[3][2] - l30 * l20 - l31 * l21) * r2;
    const double d3 = B.a[3][3] - l30 * l30 - l31 * l31 - l32 * l32, r3 = rsqrt(d3);
    bad = bad || !(B.a[0][0] > 0.0) || !(d1 > 0.0) || !(d2 > 0.0) || !(d3 > 0.0);
    const double i10 = -l10 * r0 * r1, i21 = -l21 * r1 * r2, i32 = -l32 * r2 * r3;
    const double i20 = -(l20 * r0 + l21 * i10) * r2, i31 = -(l31 * r1 + l32 * i21) * r3;
    const double i30 = -(l30 * r0 + l31 * i10 + l32 * i20) * r3;
    double2* o = sLiAll[jb];
    o[0] = make_double2(r0, 0.0); o[1] = make_double2(0.0, 0.0); o[2] = make_double2(i10, r1); o[3] = make_double2(0.0, 0.0);
    o[4] = make_double2(i20, i21); o[5] = make_double2(r2, 0.0); o[6] = make_double2(i30, i31); o[7] = make_double2(i32, r3);
  };
  for (int jb = 0; jb < nb; ++jb) {
    diag(B0, jb); diag(B1, jb);
    __syncthreads();
    rb_panel(B0, jb, nb, sLiAll[jb], sPan, ys); rb_panel(B1, jb, nb, sLiAll[jb], sPan, ys);
    __syncthreads();
    rb_update(B0, jb, sPan); rb_update(B1, jb, sPan);
  }
  if (bad) failed = 1;
  __syncthreads();
  if (failed) { if (tid == 0) st->lin_fail = 1; return; }
  // backward substitution: x_jb = Li^T (y_jb - sum_{bi > jb} L(bi,jb)^T x_bi)
  for (int jb = nb - 1; jb >= 0; --jb) {
    rb_back(B0, jb, nb, xs, part); rb_back(B1, jb, nb, xs, part);
    __syncthreads();
    if (tid < 32) {
      double v[4] = {0.0, 0.0, 0.0, 0.0};
      for (int bi = jb + 1 + tid; bi < nb; bi += 32) {
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] += part[bi][c];
      }
#pragma unroll
      for (int m = 1; m < 32; m <<= 1) {
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] += __shfl_xor_sync(0xffffffffu, v[c], m);
      }
      if (tid < 4) {   // lane m: x[m] = sum_{c >= m} Li[c][m] r[c]
        const double* Li = reinterpret_cast<const double*>(sLiAll[jb]);
        double x = 0.0;
#pragma unroll
        for (int c = 0; c < 4; ++c) if (c >= tid) x += Li[4 * c + tid] * (ys[4 * jb + c] - v[c]);
        xs[4 * jb + tid] = x;
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < n; i += kRbThreads) y[i] = xs[i];
}

// ===================================================================================================
// K7: preconditioned conjugate gradients on the block-sparse reduced system S y = rhs (3x3-block CSR, full
// symmetric storage). One cooperative persistent kernel runs every CG iteration; grid-wide barriers replace
// launches, dot products are per-CTA partial sums added in a fixed order by every CTA (no atomics).
// Preconditioner: block-Jacobi (inverse of the 3x3 diagonal blocks) = Ceres' SCHUR_JACOBI for this block size.
// SpMV: one warp per block row; lanes stride over the row's contiguous 9-double blocks (coalesced).
// ===================================================================================================
struct PcgBuffers { double* x; double* r; double* z; double* p; double* Ap; double* Minv; double* part; int max_iter; double tol; };
__device__ __forceinline__ double pcg_grid_sum(cooperative_groups::grid_group& grid, double v, double* part, int slot, double* sm) {
  const double t = block_sum<256>(v, sm);
  if (threadIdx.x == 0) part[(size_t)slot * gridDim.x + blockIdx.x] = t;
  grid.sync();
  double acc = 0.0;
  for (int i = 0; i < (int)gridDim.x; ++i) acc += part[(size_t)slot * gridDim.x + i];   // same order in every CTA
  return acc;
}
__global__ void __launch_bounds__(256) pcg_kernel(DeviceProblem P, const double* __restrict__ S, const double* __restrict__ rhs,
                                                  PcgBuffers B, LMState* st) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ double sm[256];
  if (!gate_run(st) || st->lin_fail) return;   // uniform across the grid: written only by earlier kernels
  const int n = P.nf_dim, nbr = P.nf;
  const int gtid = blockIdx.x * 256 + threadIdx.x, gsz = gridDim.x * 256;
  const int gwarp = gtid >> 5, nwarp = gsz >> 5, lane = threadIdx.x & 31;
  // block-Jacobi preconditioner: invert the 3x3 diagonal blocks (symmetric positive definite)
  int bad = 0;
  for (int i = gtid; i < nbr; i += gsz) {
    const double* D = S + (size_t)P.diag_slot[i] * 9;
    const double a = D[0], b = D[1], c = D[2], d = D[4], e = D[5], f = D[8];
    const double A0 = d * f - e * e, A1 = c * e - b * f, A2 = b * e - c * d;
    const double det = a * A0 + b * A1 + c * A2;
    if (!(det > 0.0) || !(a > 0.0)) bad = 1;
    const double id = 1.0 / det;
    double* M = B.Minv + (size_t)i * 9;
    M[0] = A0 * id; M[1] = A1 * id; M[2] = A2 * id; M[3] = A1 * id; M[4] = (a * f - c * c) * id; M[5] = (b * c - a * e) * id;
    M[6] = A2 * id; M[7] = M[5]; M[8] = (a * d - b * b) * id;
  }
  // make the failure decision uniform across the grid before anyone acts on it (st->lin_fail is only written at
  // the very end: a CTA that starts late must still see the value the launch began with)
  int fail = pcg_grid_sum(grid, (double)bad, B.part, 0, sm) > 0.0; __syncthreads();
  // x = 0, r = b, z = M^-1 r, p = z
  double bb = 0.0, rz = 0.0;
  grid.sync();
  for (int i = gtid; i < n; i += gsz) {
    const int br = i / 3, k = i % 3;
    const double* M = B.Minv + (size_t)br * 9 + 3 * k;
    const double z = M[0] * rhs[3 * br] + M[1] * rhs[3 * br + 1] + M[2] * rhs[3 * br + 2];
    B.x[i] = 0.0; B.r[i] = rhs[i]; B.z[i] = z; B.p[i] = z;
    bb += rhs[i] * rhs[i]; rz += rhs[i] * z;
  }
  bb = pcg_grid_sum(grid, bb, B.part, 0, sm); __syncthreads();
  rz = pcg_grid_sum(grid, rz, B.part, 1, sm); __syncthreads();
  const double stop = B.tol * B.tol * bb;
  int it = 0;
  if (bb > 0.0 && !fail) {
    for (it = 1; it <= B.max_iter; ++it) {
      // Ap = S p (warp per block row) and p.Ap
      double pAp = 0.0;
      for (int br = gwarp; br < nbr; br += nwarp) {
        const int e0 = P.bsr_rowptr[br] * 9, e1 = P.bsr_rowptr[br + 1] * 9;
        double y0 = 0.0, y1 = 0.0, y2 = 0.0;
        for (int e = e0 + lane; e < e1; e += 32) {
          const int blk = e / 9, w = e - 9 * blk, k = w / 3, c = w - 3 * k;
          const double v = S[e] * B.p[3 * P.bsr_col[blk] + c];
          if (k == 0) y0 += v; else if (k == 1) y1 += v; else y2 += v;
        }
#pragma unroll
        for (int m = 1; m < 32; m <<= 1) { y0 += __shfl_xor_sync(0xffffffffu, y0, m); y1 += __shfl_xor_sync(0xffffffffu, y1, m); y2 += __shfl_xor_sync(0xffffffffu, y2, m); }
        if (lane == 0) {
          B.Ap[3 * br] = y0; B.Ap[3 * br + 1] = y1; B.Ap[3 * br + 2] = y2;
          pAp += y0 * B.p[3 * br] + y1 * B.p[3 * br + 1] + y2 * B.p[3 * br + 2];
        }
      }
      pAp = pcg_grid_sum(grid, pAp, B.part, 2, sm); __syncthreads();
      if (!(pAp > 0.0)) { fail = 1; break; }   // not positive definite (uniform decision)
      const double alpha = rz / pAp;
      double rr = 0.0;
      for (int i = gtid; i < n; i += gsz) {
        B.x[i] += alpha * B.p[i];
        const double r = B.r[i] - alpha * B.Ap[i];
        B.r[i] = r; rr += r * r;
      }
      rr = pcg_grid_sum(grid, rr, B.part, 3, sm); __syncthreads();
      if (rr <= stop) break;
      double rz_new = 0.0;
      for (int i = gtid; i < n; i += gsz) {
        const int br = i / 3, k = i % 3;
        const double* M = B.Minv + (size_t)br * 9 + 3 * k;
        const double z = M[0] * B.r[3 * br] + M[1] * B.r[3 * br + 1] + M[2] * B.r[3 * br + 2];
        B.z[i] = z; rz_new += B.r[i] * z;
      }
      rz_new = pcg_grid_sum(grid, rz_new, B.part, 4, sm); __syncthreads();
      const double beta = rz_new / rz;
      rz = rz_new;
      for (int i = gtid; i < n; i += gsz) B.p[i] = B.z[i] + beta * B.p[i];
      grid.sync();
    }
  }
  if (gtid == 0) { st->pcg_iters = it > B.max_iter ? B.max_iter : it; if (fail) st->lin_fail = 1; }
}

// ---------------------------------------------------------------------------------------------------
// K7, sharded solve: the same preconditioned conjugate gradients with the BLOCK ROWS of S partitioned over the ranks
// (north star: "the PCG SpMV is partitioned by camera block"). Every rank holds the summed S (the exchange step), owns the
// block rows [rb0, rb1) and keeps x, r, z only for them; the direction p is kept whole. Per CG iteration: SpMV on the own
// rows + the own share of p.Ap, one scalar all-reduce; update of the own x, r, z, and one all-reduce that both assembles z
// (every rank contributes its slice of a zero-padded vector) and sums r.r and r.z; p = z + beta p on every rank. The
// host enqueues the iterations and looks at the (rank-identical) `done` flag every few of them. All sums are fixed-order
// inside a rank; across ranks NCCL's order applies.
struct PcgDistState { double rz, stop; int done, fail, iters; };
struct PcgDistBuf { double* x; double* r; double* p; double* Ap; double* Minv; double* zsend; double* zfull; double* part; double* scal; PcgDistState* ps;
                    int rb0, rb1, nblk; int max_iter; double tol; };
// sums of per-CTA partials: dst[k] = sum_i part[k * nblk + i], k < nk, fixed order
__global__ void __launch_bounds__(32) pcgd_sum_kernel(const double* __restrict__ part, int nblk, int nk, double* __restrict__ dst) {
  if ((int)threadIdx.x < nk) { double acc = 0.0; for (int i = 0; i < nblk; ++i) acc += part[(size_t)threadIdx.x * nblk + i]; dst[threadIdx.x] = acc; }
}
// own rows: Minv, x = 0, r = b, z = Minv r (into the zero-padded send vector); partials of b.b, r.z and the count of bad blocks
__global__ void __launch_bounds__(256) pcgd_init_kernel(DeviceProblem P, const double* __restrict__ S, const double* __restrict__ rhs, PcgDistBuf B, const LMState* st) {
  __shared__ double sm[256];
  const int br = B.rb0 + blockIdx.x * 256 + threadIdx.x;
  double bb = 0.0, rz = 0.0, bad = 0.0;
  if (br < B.rb1) {
    const double* D = S + (size_t)P.diag_slot[br] * 9;
    const double a = D[0], b = D[1], c = D[2], d = D[4], e = D[5], f = D[8];
    const double A0 = d * f - e * e, A1 = c * e - b * f, A2 = b * e - c * d;
    const double det = a * A0 + b * A1 + c * A2;
    if (!(det > 0.0) || !(a > 0.0)) bad = 1.0;
    const double id = 1.0 / det;
    double* M = B.Minv + (size_t)br * 9;
    M[0] = A0 * id; M[1] = A1 * id; M[2] = A2 * id; M[3] = A1 * id; M[4] = (a * f - c * c) * id; M[5] = (b * c - a * e) * id;
    M[6] = A2 * id; M[7] = M[5]; M[8] = (a * d - b * b) * id;
    const double r0 = rhs[3 * br], r1 = rhs[3 * br + 1], r2 = rhs[3 * br + 2];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double z = M[3 * k] * r0 + M[3 * k + 1] * r1 + M[3 * k + 2] * r2, rk = k == 0 ? r0 : k == 1 ? r1 : r2;
      B.x[3 * br + k] = 0.0; B.r[3 * br + k] = rk; B.zsend[3 * br + k] = z;
      bb += rk * rk; rz += rk * z;
    }
  }
  const double t0 = block_sum<256>(bb, sm); __syncthreads();
  const double t1 = block_sum<256>(rz, sm); __syncthreads();
  const double t2 = block_sum<256>(bad, sm);
  if (threadIdx.x == 0) { B.part[blockIdx.x] = t0; B.part[B.nblk + blockIdx.x] = t1; B.part[2 * B.nblk + blockIdx.x] = t2; }
  if (blockIdx.x == 0 && threadIdx.x == 0) { B.ps->done = (!gate_run(st) || st->lin_fail) ? 1 : 0; B.ps->fail = 0; B.ps->iters = 0; }
}
// every rank, whole vector: p = z; the scalars of the start
__global__ void __launch_bounds__(256) pcgd_start_kernel(int n, PcgDistBuf B) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) B.p[i] = B.zfull[i];
}
__global__ void pcgd_start_ctrl_kernel(int n, PcgDistBuf B) {
  PcgDistState* ps = B.ps;
  const double bb = B.zfull[n], rz = B.zfull[n + 1], bad = B.zfull[n + 2];
  ps->rz = rz; ps->stop = B.tol * B.tol * bb;
  if (bad > 0.0) { ps->fail = 1; ps->done = 1; }
  if (!(bb > 0.0)) ps->done = 1;
}
// own rows: Ap = S p (warp per block row), per-CTA partial of p.Ap
__global__ void __launch_bounds__(256) pcgd_spmv_kernel(DeviceProblem P, const double* __restrict__ S, PcgDistBuf B) {
  __shared__ double sm[256];
  double pAp = 0.0;
  if (!B.ps->done) {
    const int br = B.rb0 + blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (br < B.rb1) {
      const int e0 = P.bsr_rowptr[br] * 9, e1 = P.bsr_rowptr[br + 1] * 9;
      double y0 = 0.0, y1 = 0.0, y2 = 0.0;
      for (int e = e0 + lane; e < e1; e += 32) {
        const int blk = e / 9, w = e - 9 * blk, k = w / 3, c = w - 3 * k;
        const double v = S[e] * B.p[3 * P.bsr_col[blk] + c];
        if (k == 0) y0 += v; else if (k == 1) y1 += v; else y2 += v;
      }
#pragma unroll
      for (int m = 1; m < 32; m <<= 1) { y0 += __shfl_xor_sync(0xffffffffu, y0, m); y1 += __shfl_xor_sync(0xffffffffu, y1, m); y2 += __shfl_xor_sync(0xffffffffu, y2, m); }
      if (lane == 0) {
        B.Ap[3 * br] = y0; B.Ap[3 * br + 1] = y1; B.Ap[3 * br + 2] = y2;
        pAp = y0 * B.p[3 * br] + y1 * B.p[3 * br + 1] + y2 * B.p[3 * br + 2];
      }
    }
  }
  const double t = block_sum<256>(pAp, sm);
  if (threadIdx.x == 0) B.part[blockIdx.x] = t;
}
// own rows: x += alpha p, r -= alpha Ap, z = Minv r; per-CTA partials of r.r and r.z. scal[0] = p.Ap summed over the ranks
__global__ void __launch_bounds__(256) pcgd_update_kernel(PcgDistBuf B) {
  __shared__ double sm[256];
  const int br = B.rb0 + blockIdx.x * 256 + threadIdx.x;
  double rr = 0.0, rz = 0.0;
  const double pAp = B.scal[0];
  if (!B.ps->done && pAp > 0.0 && br < B.rb1) {
    const double alpha = B.ps->rz / pAp;
    double r[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { B.x[3 * br + k] += alpha * B.p[3 * br + k]; r[k] = B.r[3 * br + k] - alpha * B.Ap[3 * br + k]; B.r[3 * br + k] = r[k]; rr += r[k] * r[k]; }
    const double* M = B.Minv + (size_t)br * 9;
#pragma unroll
    for (int k = 0; k < 3; ++k) { const double z = M[3 * k] * r[0] + M[3 * k + 1] * r[1] + M[3 * k + 2] * r[2]; B.zsend[3 * br + k] = z; rz += r[k] * z; }
  }
  const double t0 = block_sum<256>(rr, sm); __syncthreads();
  const double t1 = block_sum<256>(rz, sm);
  if (threadIdx.x == 0) { B.part[blockIdx.x] = t0; B.part[B.nblk + blockIdx.x] = t1; }
}
// every rank, whole vector: p = z + beta p unless the iteration ended the solve
__global__ void __launch_bounds__(256) pcgd_dir_kernel(int n, PcgDistBuf B) {
  const double pAp = B.scal[0], rr = B.zfull[n], rzn = B.zfull[n + 1];
  if (B.ps->done || !(pAp > 0.0) || rr <= B.ps->stop) return;
  const double beta = rzn / B.ps->rz;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) B.p[i] = B.zfull[i] + beta * B.p[i];
}
__global__ void pcgd_ctrl_kernel(int n, PcgDistBuf B) {
  PcgDistState* ps = B.ps;
  if (ps->done) return;
  const double pAp = B.scal[0], rr = B.zfull[n], rzn = B.zfull[n + 1];
  ps->iters++;
  if (!(pAp > 0.0)) { ps->fail = 1; ps->done = 1; }          // not positive definite
  else if (rr <= ps->stop || ps->iters >= B.max_iter) ps->done = 1;
  else ps->rz = rzn;
}
__global__ void pcgd_finish_kernel(PcgDistBuf B, LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  st->pcg_iters = B.ps->iters;
  if (B.ps->fail) st->lin_fail = 1;
}

// ===================================================================================================
// K5c: back-substitution of the landmark steps and the model cost change
//   y_e = Minv (g_e - sum E^T F y_f);  model_cost_change = sum_rows m (r - m/2), m = F y_f + E y_e  (delta = -y)
// ===================================================================================================
template <int G>
__global__ void __launch_bounds__(256) backsub_kernel(DeviceProblem P, const double* __restrict__ Jrec,
                                                      const double* __restrict__ lm_minv, const double* __restrict__ grad,
                                                      const double* __restrict__ yf, double* __restrict__ y,
                                                      double* __restrict__ model_part, const LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  __shared__ double sm[256];
  const int gtid = blockIdx.x * 256 + threadIdx.x;
  const int l = gtid / G, sub = gtid % G;
  const bool valid = l < P.ne;
  const int start = valid ? P.lm_ptr[l] : 0, end = valid ? P.lm_ptr[l + 1] : 0;
  // The first kBsKeep observations of a lane stay in registers between the two passes (E, r and u = F y_pose): with four
  // lanes per landmark that covers tracks of up to 12, so the records are read from HBM once. They are fetched as in
  // landmark_kernel: the warp loads 16-byte pieces that are consecutive in memory and every lane takes its own record
  // from a per-warp shared-memory buffer. Longer tracks re-read the rest.
  constexpr int kBsKeep = 3;
  __shared__ double2 sRec[8][32 * 9];
  const int lane = threadIdx.x & 31;
  double2* const buf = sRec[threadIdx.x >> 5];
  double2 ke0[kBsKeep], ke1[kBsKeep], ke2[kBsKeep], kr[kBsKeep]; double ku0[kBsKeep], ku1[kBsKeep];
#pragma unroll
  for (int b = 0; b < kBsKeep; ++b) {
    const int o = start + sub + b * G;
    const int pm = o < end ? P.obs_pm[o] : -1;
    int tq = -1, tt = -1;
    if (pm >= 0) { const int pose = P.obs_pose[o]; tq = P.pose_q_tan[pose]; tt = P.pose_t_tan[pose]; }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int q = lane + 32 * j, slot = q >> 3, piece = q & 7;
      const int spm = __shfl_sync(0xffffffffu, pm, slot);
      buf[slot * 9 + piece] = spm >= 0 ? reinterpret_cast<const double2*>(Jrec + (size_t)spm * kJRec)[piece] : make_double2(0.0, 0.0);
    }
    __syncwarp();
    double yp[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) { yp[k] = tq >= 0 ? yf[tq - P.ne_dim + k] : 0.0; yp[3 + k] = tt >= 0 ? yf[tt - P.ne_dim + k] : 0.0; }
    const double2* rc = buf + lane * 9;
    const double2 e0 = rc[0], e1 = rc[1], e2 = rc[2], f0 = rc[4], f1 = rc[5], f2 = rc[6]; const double ts = rc[7].x;
    ke0[b] = e0; ke1[b] = e1; ke2[b] = e2; kr[b] = rc[3];
    __syncwarp();
    ku0[b] = f0.x * yp[0] + f0.y * yp[1] + f1.x * yp[2] + ts * (e0.x * yp[3] + e0.y * yp[4] + e1.x * yp[5]);
    ku1[b] = f1.y * yp[0] + f2.x * yp[1] + f2.y * yp[2] + ts * (e1.y * yp[3] + e2.x * yp[4] + e2.y * yp[5]);
  }
  double t0 = 0, t1 = 0, t2 = 0;
#pragma unroll
  for (int b = 0; b < kBsKeep; ++b) { t0 += ke0[b].x * ku0[b] + ke1[b].y * ku1[b]; t1 += ke0[b].y * ku0[b] + ke2[b].x * ku1[b]; t2 += ke1[b].x * ku0[b] + ke2[b].y * ku1[b]; }
  for (int o = start + sub + kBsKeep * G; o < end; o += G) {
    const double2* rec = reinterpret_cast<const double2*>(Jrec + (size_t)P.obs_pm[o] * kJRec);
    const int pose = P.obs_pose[o];
    const int tq = P.pose_q_tan[pose], tt = P.pose_t_tan[pose];
    double yp[6];
    for (int k = 0; k < 3; ++k) { yp[k] = tq >= 0 ? yf[tq - P.ne_dim + k] : 0.0; yp[3 + k] = tt >= 0 ? yf[tt - P.ne_dim + k] : 0.0; }
    const double2 f0 = rec[4], f1 = rec[5], f2 = rec[6]; const double ts = rec[7].x;
    const double2 e0 = rec[0], e1 = rec[1], e2 = rec[2];
    const double u0 = f0.x * yp[0] + f0.y * yp[1] + f1.x * yp[2] + ts * (e0.x * yp[3] + e0.y * yp[4] + e1.x * yp[5]);
    const double u1 = f1.y * yp[0] + f2.x * yp[1] + f2.y * yp[2] + ts * (e1.y * yp[3] + e2.x * yp[4] + e2.y * yp[5]);
    t0 += e0.x * u0 + e1.y * u1; t1 += e0.y * u0 + e2.x * u1; t2 += e1.x * u0 + e2.y * u1;
  }
#pragma unroll
  for (int m = 1; m < G; m <<= 1) {
    t0 += __shfl_xor_sync(0xffffffffu, t0, m); t1 += __shfl_xor_sync(0xffffffffu, t1, m); t2 += __shfl_xor_sync(0xffffffffu, t2, m);
  }
  double ye0 = 0, ye1 = 0, ye2 = 0;
  if (valid) {
    const double* mi = lm_minv + (size_t)l * 6;
    const double b0 = grad[3 * l] - t0, b1 = grad[3 * l + 1] - t1, b2 = grad[3 * l + 2] - t2;
    ye0 = mi[0] * b0 + mi[1] * b1 + mi[2] * b2; ye1 = mi[1] * b0 + mi[3] * b1 + mi[4] * b2; ye2 = mi[2] * b0 + mi[4] * b1 + mi[5] * b2;
    if (sub == 0) { y[3 * l] = ye0; y[3 * l + 1] = ye1; y[3 * l + 2] = ye2; }
  }
  double mc = 0.0;
#pragma unroll
  for (int b = 0; b < kBsKeep; ++b) {   // empty slots hold zeros and add nothing
    const double m0 = ku0[b] + ke0[b].x * ye0 + ke0[b].y * ye1 + ke1[b].x * ye2;
    const double m1 = ku1[b] + ke1[b].y * ye0 + ke2[b].x * ye1 + ke2[b].y * ye2;
    mc += m0 * (kr[b].x - 0.5 * m0) + m1 * (kr[b].y - 0.5 * m1);
  }
  for (int o = start + sub + kBsKeep * G; o < end; o += G) {
    const double2* rec = reinterpret_cast<const double2*>(Jrec + (size_t)P.obs_pm[o] * kJRec);
    const int pose = P.obs_pose[o];
    const int tq = P.pose_q_tan[pose], tt = P.pose_t_tan[pose];
    double yp[6];
    for (int k = 0; k < 3; ++k) { yp[k] = tq >= 0 ? yf[tq - P.ne_dim + k] : 0.0; yp[3 + k] = tt >= 0 ? yf[tt - P.ne_dim + k] : 0.0; }
    const double2 f0 = rec[4], f1 = rec[5], f2 = rec[6]; const double ts = rec[7].x;
    const double2 e0 = rec[0], e1 = rec[1], e2 = rec[2], r = rec[3];
    const double m0 = f0.x * yp[0] + f0.y * yp[1] + f1.x * yp[2] + ts * (e0.x * yp[3] + e0.y * yp[4] + e1.x * yp[5]) + e0.x * ye0 + e0.y * ye1 + e1.x * ye2;
    const double m1 = f1.y * yp[0] + f2.x * yp[1] + f2.y * yp[2] + ts * (e1.y * yp[3] + e2.x * yp[4] + e2.y * yp[5]) + e1.y * ye0 + e2.x * ye1 + e2.y * ye2;
    mc += m0 * (r.x - 0.5 * m0) + m1 * (r.y - 0.5 * m1);
  }
  const double tot = block_sum<256>(mc, sm);
  if (threadIdx.x == 0) model_part[blockIdx.x] = tot;
}
// model cost change of the generic factors + copy y_f into the full step vector. One warp per factor: lane c holds the
// step entry of tangent column c (<= 30 columns), rows are reduced with a butterfly.
__global__ void __launch_bounds__(128) gf_model_kernel(DeviceProblem P, const double* __restrict__ gfJ,
                                                       const double* __restrict__ gfR, const double* __restrict__ yf,
                                                       double* __restrict__ y, double* __restrict__ model_part, const LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  const int tid = blockIdx.x * 128 + threadIdx.x;
  for (int j = tid; j < P.nf_dim; j += gridDim.x * 128) y[P.ne_dim + j] = yf[j];
  const int f = tid >> 5, lane = threadIdx.x & 31;
  if (f >= P.num_gf) return;
  const int nb = P.gf_nblk[f], nr = P.gf_nres[f], ncol = 3 * nb;
  const double* J = gfJ + P.gf_joff[f]; const double* R = gfR + P.gf_roff[f];
  double yl = 0.0;
  if (lane < ncol) { const int t = P.blk_tan[P.gf_blocks[f * kGfMaxBlocks + lane / 3]]; yl = t >= 0 ? yf[t - P.ne_dim + lane % 3] : 0.0; }
  double mc = 0.0;
  for (int i = 0; i < nr; ++i) {
    double m = lane < ncol ? J[i * ncol + lane] * yl : 0.0;
#pragma unroll
    for (int w = 1; w < 32; w <<= 1) m += __shfl_xor_sync(0xffffffffu, m, w);
    mc += m * (R[i] - 0.5 * m);
  }
  if (lane == 0) model_part[f] = mc;
}

// candidate point x_cand = x (+) (-y) and the ambient norms ParameterToleranceReached needs.
__global__ void __launch_bounds__(256) candidate_kernel(DeviceProblem P, const double* __restrict__ x,
                                                        const double* __restrict__ y, double* __restrict__ xc,
                                                        double* __restrict__ part_step, double* __restrict__ part_cand,
                                                        const LMState* st) {
  if (!gate_run(st) || st->lin_fail) return;
  __shared__ double sm[256];
  const int i = blockIdx.x * 256 + threadIdx.x;
  double ss = 0.0, cs = 0.0;
  if (i < P.num_active && P.active_blocks[i] >= 0) {
    const int b = P.active_blocks[i], off = P.blk_off[b], t = P.blk_tan[b];
    if (P.blk_manifold[b] == BSNLS_MANIFOLD_QUATERNION) {
      const double d[3] = {-y[t], -y[t + 1], -y[t + 2]};
      double out[4]; quat_plus(x + off, d, out);
      for (int k = 0; k < 4; ++k) { xc[off + k] = out[k]; const double dd = x[off + k] - out[k]; ss += dd * dd; cs += out[k] * out[k]; }
    } else {
      for (int k = 0; k < P.blk_size[b]; ++k) { const double v = x[off + k] + (-y[t + k]); xc[off + k] = v; const double dd = x[off + k] - v; ss += dd * dd; cs += v * v; }
    }
  }
  const double a = block_sum<256>(ss, sm); __syncthreads();
  const double c = block_sum<256>(cs, sm);
  if (threadIdx.x == 0) { part_step[blockIdx.x] = a; part_cand[blockIdx.x] = c; }
}
__global__ void __launch_bounds__(256) xnorm_kernel(DeviceProblem P, const double* __restrict__ x, double* __restrict__ part) {
  __shared__ double sm[256];
  const int i = blockIdx.x * 256 + threadIdx.x;
  double s = 0.0;
  if (i < P.num_active && P.active_blocks[i] >= 0) { const int b = P.active_blocks[i], off = P.blk_off[b]; for (int k = 0; k < P.blk_size[b]; ++k) s += x[off + k] * x[off + k]; }
  const double a = block_sum<256>(s, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = a;
}

// Deterministic final reductions: one CTA walks a list of (array, length, op) descriptors.
// red[] layout: [0, RED_NUM) sums over rank-local data (observations, landmarks) — all-reduced across ranks in the
// sharded solve; [RED_NUM, 2 RED_NUM) sums over replicated data (generic factors, f-blocks); then local max, replicated max.
constexpr int kRedLocalMax = 2 * RED_NUM, kRedReplMax = 2 * RED_NUM + 1, kRedSize = 2 * RED_NUM + 2;
__device__ __forceinline__ double red_sum(const double* red, int slot) { return red[slot] + red[RED_NUM + slot]; }
__device__ __forceinline__ double red_max(const double* red) { return fmax(red[kRedLocalMax], red[kRedReplMax]); }
struct RedDesc { const double* src; int n; int op; int slot; };  // op 0 = sum, 1 = max; slot already offset by RED_NUM when replicated
struct RedList { RedDesc d[10]; int count; int gate; int keep; };   // gate 0 always, 1 run, 2 jac; keep: add to red[] instead of starting from zero
__global__ void __launch_bounds__(256) reduce_kernel(RedList L, double* __restrict__ red, const LMState* st) {
  if (L.gate == 1 && !gate_run(st)) return;
  if (L.gate == 2 && !gate_jac(st)) return;
  __shared__ double sm[256];
  if (!L.keep && threadIdx.x < kRedSize) red[threadIdx.x] = 0.0;   // slots not produced by this phase must not carry stale sums
  __syncthreads();
  for (int q = 0; q < L.count; ++q) {
    const RedDesc d = L.d[q];
    double v = 0.0;
    for (int i = threadIdx.x; i < d.n; i += 256) v = d.op ? fmax(v, d.src[i]) : v + d.src[i];
    const double t = d.op ? block_max<256>(v, sm) : block_sum<256>(v, sm);
    if (threadIdx.x == 0) red[d.slot] = (q > 0 && L.d[q - 1].slot == d.slot) ? red[d.slot] + t : t;
    __syncthreads();
  }
}

// ===================================================================================================
// K9: device-side LM controller (ceres TrustRegionMinimizer + LevenbergMarquardtStrategy).
// ===================================================================================================
__device__ __forceinline__ void push_record(LMState* st, bsnls_iteration* recs, const bsnls_iteration& r) {
  if (st->num_records < st->max_records) recs[st->num_records] = r;
  st->num_records++;
}
// FinalizeIterationAndCheckIfMinimizerCanContinue after a record was pushed
__device__ __forceinline__ void finalize_checks(LMState* st, const bsnls_iteration& r, const Options& opt) {
  if (r.step_is_successful) st->num_successful++;
  if (r.iteration >= opt.max_num_iterations) { st->done = 1; st->termination = BSNLS_NO_CONVERGENCE; st->message_id = MSG_MAX_ITER; return; }
  if (r.step_is_successful && r.gradient_max_norm <= opt.gradient_tolerance) { st->done = 1; st->termination = BSNLS_CONVERGENCE; st->message_id = MSG_GRAD_TOL; return; }
  if (st->radius < opt.min_radius) { st->done = 1; st->termination = BSNLS_CONVERGENCE; st->message_id = MSG_MIN_RADIUS; return; }
}
// after a Jacobian evaluation: IterationZero, or the tail of HandleSuccessfulStep
__global__ void ctrl_after_jacobian_kernel(LMState* st, const double* red, bsnls_iteration* recs, Options opt) {
  if (!gate_jac(st)) return;
  st->x_cost = red_sum(red, RED_COST); st->fixed_cost = red_sum(red, RED_FIXED); st->accepted = 0;
  st->grad_max = red_max(red); st->grad_norm = sqrt(red_sum(red, RED_GRAD_SQ));
  bsnls_iteration r;
  r.iteration = st->iteration; r.step_is_valid = 1; r.step_is_successful = 1;
  r.linear_solver_iterations = st->iteration == 0 ? 0 : st->pend_linear_iters;
  r.cost = st->x_cost + st->fixed_cost;
  r.cost_change = st->iteration == 0 ? 0.0 : st->pend_cost_change;
  r.gradient_max_norm = st->grad_max; r.gradient_norm = st->grad_norm;
  r.step_norm = st->iteration == 0 ? 0.0 : st->pend_step_norm;
  r.relative_decrease = st->iteration == 0 ? 0.0 : st->pend_relative_decrease;
  r.trust_region_radius = st->radius;
  st->need_jacobian = 0; st->scale_ready = 1;
  if (!isfinite(st->x_cost)) { st->done = 1; st->termination = BSNLS_FAILURE; st->message_id = MSG_NONFINITE; push_record(st, recs, r); return; }
  push_record(st, recs, r);
  finalize_checks(st, r, opt);
}
// end of an iteration: step validity, tolerances, acceptance, trust-region update
__global__ void ctrl_end_iteration_kernel(LMState* st, const double* red, bsnls_iteration* recs, Options opt) {
  if (!gate_run(st)) return;
  st->iteration += 1;
  bsnls_iteration r;
  r.iteration = st->iteration; r.step_is_valid = 0; r.step_is_successful = 0; r.linear_solver_iterations = st->pcg_iters;
  r.cost = st->x_cost + st->fixed_cost; r.cost_change = 0.0; r.gradient_max_norm = st->grad_max; r.gradient_norm = st->grad_norm;
  r.step_norm = 0.0; r.relative_decrease = 0.0;
  const double mc = red_sum(red, RED_MODEL);
  const bool solved = !st->lin_fail && isfinite(mc) && isfinite(red_sum(red, RED_STEP_SQ));
  st->lin_fail = 0;
  if (!solved || !(mc > 0.0)) {  // HandleInvalidStep
    st->num_invalid += 1;
    if (st->num_invalid >= opt.max_invalid) { st->done = 1; st->termination = BSNLS_FAILURE; st->message_id = MSG_INVALID_STEPS; return; }
    st->radius = st->radius / st->decrease_factor; st->decrease_factor *= 2.0;
    st->num_unsuccessful++;
    r.trust_region_radius = st->radius;
    push_record(st, recs, r); finalize_checks(st, r, opt);
    return;
  }
  st->num_invalid = 0;
  r.step_is_valid = 1;
  st->model_cost_change = mc;
  st->step_norm = sqrt(red_sum(red, RED_STEP_SQ)); st->cand_norm = sqrt(red_sum(red, RED_CAND_SQ)); st->cand_cost = red_sum(red, RED_CAND_COST);
  r.step_norm = st->step_norm;
  // ParameterToleranceReached
  if (st->step_norm <= opt.parameter_tolerance * (st->x_norm + opt.parameter_tolerance)) {
    st->done = 1; st->termination = BSNLS_CONVERGENCE; st->message_id = MSG_PARAM_TOL; return; }
  // FunctionToleranceReached
  r.cost_change = st->x_cost - st->cand_cost;
  if (fabs(r.cost_change) <= opt.function_tolerance * st->x_cost) {
    st->done = 1; st->termination = BSNLS_CONVERGENCE; st->message_id = MSG_FUNC_TOL; return; }
  r.relative_decrease = (st->x_cost - st->cand_cost) / mc;
  if (r.relative_decrease > opt.min_relative_decrease) {  // HandleSuccessfulStep (completed by ctrl_after_jacobian)
    st->accepted = 1; st->need_jacobian = 1;
    st->x_norm = st->cand_norm;
    st->pend_cost_change = r.cost_change; st->pend_step_norm = r.step_norm; st->pend_relative_decrease = r.relative_decrease;
    st->pend_linear_iters = st->pcg_iters;
    const double q = 2.0 * r.relative_decrease - 1.0;
    st->radius = st->radius / fmax(1.0 / 3.0, 1.0 - q * q * q);
    st->radius = fmin(opt.max_radius, st->radius);
    st->decrease_factor = 2.0;
  } else {  // HandleUnsuccessfulStep
    st->radius = st->radius / st->decrease_factor; st->decrease_factor *= 2.0;
    st->num_unsuccessful++;
    r.cost = st->cand_cost + st->fixed_cost; r.trust_region_radius = st->radius;
    push_record(st, recs, r); finalize_checks(st, r, opt);
  }
}
__global__ void __launch_bounds__(256) apply_step_kernel(double* __restrict__ x, const double* __restrict__ xc, int n, LMState* st) {
  if (!st->accepted) return;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) x[i] = xc[i];
}
__global__ void init_xnorm_kernel(LMState* st, const double* red) { st->x_norm = sqrt(red_sum(red, RED_X_SQ)); }
// complete parameter vector on every rank of the sharded solve: each entry is kept by its owner, then summed
__global__ void __launch_bounds__(256) mask_owned_kernel(double* __restrict__ x, const unsigned char* __restrict__ owned, int n) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n && !owned[i]) x[i] = 0.0;
}
// right-hand side e_j (covariance columns: S^-1 e_j)
__global__ void __launch_bounds__(256) unit_rhs_kernel(double* __restrict__ rhs, int n, int j) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) rhs[i] = i == j ? 1.0 : 0.0;
}
__global__ void set_done_kernel(LMState* st, int termination, int msg) { if (!st->done) { st->done = 1; st->termination = termination; st->message_id = msg; } }

}  // namespace bsnls
