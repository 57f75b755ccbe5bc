/*
 * bsnls.h — C-ABI of the B200-native nonlinear least-squares back end for beam_slam.
 *
 * This is the drop-in boundary for ONE path of the reference: the Levenberg-Marquardt solve that
 * fuse_graphs::HashGraph::optimize hands to Ceres
 *   (/root/reference/dependencies/fuse/fuse_graphs/src/hash_graph.cpp:416-426, createProblem :475-522),
 * called by bs_optimizers::FixedLagSmoother (bs_optimizers/src/fixed_lag_smoother.cpp:281) and by the
 * global-mapper refinement (bs_models/src/lib/global_mapping/submap_refinement.cpp:114 ...).
 *
 * Plain C: pointers, sizes and PODs only. The caller owns every host array passed in; the library copies
 * what it needs in bsnls_set_problem and owns all device memory. All arithmetic is IEEE double.
 * Every function returns 0 on success and a negative bsnls_status on error; bsnls_last_error() gives text.
 * Nothing here falls back to a CPU implementation: without a CUDA device every call fails with
 * BSNLS_ERR_CUDA.
 */
#ifndef BSNLS_H_
#define BSNLS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum bsnls_status {
  BSNLS_OK = 0,
  BSNLS_ERR_INVALID = -1,     /* bad argument / invalid options (ceres_options.cpp:195-199 throws here) */
  BSNLS_ERR_UNSUPPORTED = -2, /* a graph feature the GPU path does not implement (bounds, unknown factor) */
  BSNLS_ERR_CUDA = -3,        /* CUDA / NCCL runtime failure, or no device */
  BSNLS_ERR_STATE = -4        /* call order violated (solve before set_problem ...) */
} bsnls_status;

/* ---- parameter blocks: one per fuse_core::Variable, in createProblem order (hash_graph.cpp:478) ---- */
enum { BSNLS_MANIFOLD_EUCLIDEAN = 0,
       /* fuse_variables::Orientation3DLocalParameterization (orientation_3d_stamped.h:91-115):
          w,x,y,z storage, x (+) d = x * AngleAxisToQuaternion(d), tangent size 3 */
       BSNLS_MANIFOLD_QUATERNION = 1 };

/* ---- residual blocks: one group per constraint type, each in createProblem order (hash_graph.cpp:507) ---- */
typedef enum bsnls_factor_type {
  /* bs_constraints::EuclideanReprojectionConstraint -> EuclideanReprojection SizedCostFunction<2,4,3,3>
     (bs_constraints/include/bs_constraints/visual/euclidean_reprojection_function.h:66-172).
     blocks: q_WORLD_BASELINK[4], t_WORLD_BASELINK[3], P_WORLD[3]; consts: pixel[2]; aux: camera index. */
  BSNLS_FACTOR_REPROJECTION = 0,
  /* bs_constraints::RelativeImuState3DStampedConstraint -> NormalDeltaImuState3DCostFunctor
     (bs_constraints/include/bs_constraints/inertial/normal_delta_imu_state_3d_cost_functor.h:59-141).
     blocks: q_i,p_i,v_i,bg_i,ba_i,q_j,p_j,v_j,bg_j,ba_j; consts (BSNLS_IMU_DELTA_STRIDE):
     dt, dq[4] (w,x,y,z), dp[3], dv[3], dq_dbg[9], dp_dbg[9], dp_dba[9], dv_dbg[9], dv_dba[9] (row-major),
     bg0[3], ba0[3], A[225] (row-major, = info_weight * sqrt_inv_cov). */
  BSNLS_FACTOR_IMU_DELTA = 1,
  /* bs_constraints::AbsoluteImuState3DStampedConstraint -> NormalPriorImuState3DCostFunctor
     (bs_constraints/include/bs_constraints/inertial/normal_prior_imu_state_3d_cost_functor.h:54-90);
     the pseudo-marginalisation prior of fixed_lag_smoother.cpp:244-268.
     blocks: q,p,v,bg,ba; consts: mean[16] (q w,x,y,z; p; v; bg; ba), A[225] row-major. */
  BSNLS_FACTOR_IMU_PRIOR = 2,
  /* fuse_constraints::RelativePose3DStampedConstraint -> NormalDeltaPose3DCostFunctor
     (fuse_constraints/include/fuse_constraints/normal_delta_pose_3d_cost_functor.h:107-148).
     blocks: p1,q1,p2,q2; consts: delta[7] (x,y,z,qw,qx,qy,qz), A[36] row-major. */
  BSNLS_FACTOR_RELATIVE_POSE = 3,
  /* bs_constraints::RelativePose3DStampedWithExtrinsicsConstraint -> DeltaPose3DWithExtrinsicsCostFunctor
     (bs_constraints/include/bs_constraints/relative_pose/delta_pose_3d_with_extrinsics_cost_functor.h:65-109).
     blocks: p1,q1,p2,q2,p_ext,q_ext; consts as RELATIVE_POSE. */
  BSNLS_FACTOR_RELATIVE_POSE_EXTRINSICS = 4,
  /* fuse_constraints::AbsolutePose3DStampedConstraint -> NormalPriorPose3DCostFunctor
     (fuse_constraints/include/fuse_constraints/normal_prior_pose_3d_cost_functor.h:86).
     blocks: p,q; consts: mean[7], A[36]. */
  BSNLS_FACTOR_ABSOLUTE_POSE = 5,
  /* fuse_constraints::RelativeOrientation3DStampedConstraint -> NormalDeltaOrientation3DCostFunctor
     (fuse_constraints/include/fuse_constraints/normal_delta_orientation_3d_cost_functor.h:92-123).
     blocks: q1,q2; consts: delta[4], A[9]. (Golden vector of test_marginalize_variables.cpp:298-356.) */
  BSNLS_FACTOR_RELATIVE_ORIENTATION = 6,
  /* bs_constraints::AbsoluteConstraint<T> / RelativeConstraint<T> on 3-vectors
     (bs_constraints/include/bs_constraints/global/absolute_constraint.h:14-28; the zero-motion bundle of
     bs_common/src/bs_common/utils.cpp:88-158). blocks: x  |  x1,x2; consts: mean|delta[3], A[9]. */
  BSNLS_FACTOR_ABSOLUTE_VEC3 = 7,
  BSNLS_FACTOR_RELATIVE_VEC3 = 8,
  /* fuse_constraints::MarginalConstraint -> MarginalCostFunction (fuse_constraints/src/marginal_cost_function.cpp:41-113):
     the linear prior that marginalizeVariables leaves on the remaining variables (marginalize_variables.cpp:166-258),
     cost(x) = || A1 (x1 (-) x1_bar) + ... + An (xn (-) xn_bar) + b ||^2, (-) = the variable's Minus (quaternion: angle-axis of
     x_bar^-1 * x, orientation_3d_stamped.h:117-128; vectors: difference); the Jacobian is Ai times the Minus Jacobian at x
     (:130-142), as the reference takes it. The only type with a variable shape: BSNLS_MARGINAL_MAX_BLOCKS block slots per
     factor in `blocks` (unused slots -1, used ones first), at most BSNLS_MARGINAL_MAX_ROWS residual rows (unused rows zero).
     consts (BSNLS_MARGINAL_STRIDE): n (number of blocks), size[10] (ambient size of block i: 3 or 4; 0 unused), b[15],
     x_bar[32] (ambient, blocks packed in order), A[15][30] row-major (three tangent columns per block, blocks in order). */
  BSNLS_FACTOR_MARGINAL = 9,
  /* bs_constraints::GravityAlignmentStampedConstraint -> GravityAlignmentCostFunctor
     (bs_constraints/include/bs_constraints/global/gravity_alignment_cost_functor.h:32-75): the x and y components of
     R(q_WORLD_BASELINK) * gravity_in_baselink, weighted. blocks: q; consts: gravity_in_baselink[3], A[4] (2x2 row-major). */
  BSNLS_FACTOR_GRAVITY_ALIGNMENT = 10,
  /* fuse_constraints::AbsoluteOrientation3DStampedConstraint -> NormalPriorOrientation3DCostFunctor
     (fuse_constraints/include/fuse_constraints/normal_prior_orientation_3d_cost_functor.h:92-125).
     blocks: q; consts: mean[4] (w,x,y,z), A[9]. (The prior of test_marginalize_variables.cpp:416-560.) */
  BSNLS_FACTOR_ABSOLUTE_ORIENTATION = 11,
  BSNLS_NUM_FACTOR_TYPES = 12
} bsnls_factor_type;

enum { BSNLS_IMU_DELTA_STRIDE = 288, BSNLS_IMU_PRIOR_STRIDE = 241, BSNLS_POSE_STRIDE = 43,
       BSNLS_ORIENTATION_STRIDE = 13, BSNLS_VEC3_STRIDE = 12, BSNLS_REPROJECTION_STRIDE = 2,
       BSNLS_GRAVITY_STRIDE = 7,
       BSNLS_MARGINAL_MAX_BLOCKS = 10, BSNLS_MARGINAL_MAX_ROWS = 15, BSNLS_MARGINAL_STRIDE = 1 + 10 + 15 + 32 + 15 * 30 };

/* fuse_loss plugins (dependencies/fuse/fuse_loss/src/ *.cpp) wrap ceres::LossFunction classes, one scale parameter `a` each.
   CAUCHY .. TUKEY are Ceres' own classes (loss_function.cc, third-party; Tukey as in Ceres 2.x); DCS, FAIR, GEMAN_MCCLURE
   and WELSCH are defined by fuse itself (fuse_loss/src/loss_function.cpp:44-103). Not covered: TolerantLoss (two
   parameters), ComposedLoss, ScaledLoss. */
enum { BSNLS_LOSS_NONE = 0, BSNLS_LOSS_CAUCHY = 1, BSNLS_LOSS_HUBER = 2, BSNLS_LOSS_SOFTLONE = 3, BSNLS_LOSS_ARCTAN = 4,
       BSNLS_LOSS_TUKEY = 5, BSNLS_LOSS_DCS = 6, BSNLS_LOSS_FAIR = 7, BSNLS_LOSS_GEMAN_MCCLURE = 8, BSNLS_LOSS_WELSCH = 9,
       BSNLS_NUM_LOSS_TYPES = 10 };

/* One camera "class" shared by many reprojection factors: the members of
   EuclideanReprojectionConstraint (euclidean_reprojection_constraint.cpp:17-30) that do not vary per
   observation. BSNLS_CAMERA_STRIDE doubles each:
   fx, fy, cx, cy, T_cam_baselink[12] (3x4 row-major), w (sqrt_information = w*I2), loss_type, loss_a, 0 */
enum { BSNLS_CAMERA_STRIDE = 20 };

typedef struct bsnls_factor_group {
  int32_t type;              /* bsnls_factor_type */
  int32_t count;
  const int32_t* blocks;     /* count x nblocks(type) parameter-block indices, Constraint::variables() order */
  const double* consts;      /* count x stride(type) */
  const int32_t* aux;        /* reprojection: camera index; otherwise NULL */
  const int32_t* loss_type;  /* count, or NULL (= none). Ignored for reprojection (loss lives in the camera) */
  const double* loss_param;  /* count, or NULL */
  const int32_t* res_index;  /* count: position of each factor in createProblem's residual-block order */
} bsnls_factor_group;

typedef struct bsnls_problem {
  int32_t num_blocks;
  const int32_t* block_size;     /* ambient size: Variable::size() */
  const int32_t* block_manifold; /* BSNLS_MANIFOLD_* : Variable::localParameterization() */
  const uint8_t* block_constant; /* Variable::holdConstant() || HashGraph::variables_on_hold_ */
  const int32_t* block_group;    /* elimination group: 0 = eliminated first (landmarks), 1 = reduced system */
  const double* block_values;    /* concatenated Variable::data(), offsets = prefix sums of block_size */
  int32_t num_cameras;
  const double* cameras;         /* num_cameras x BSNLS_CAMERA_STRIDE */
  int32_t num_groups;
  const bsnls_factor_group* groups;
} bsnls_problem;

/* Subset of ceres::Solver::Options that fuse loads (fuse_core/src/ceres_options.cpp:75-200) and that the
   trust-region path consults. Defaults = Ceres defaults (bsnls_default_options). */
enum { BSNLS_LINEAR_DENSE_SCHUR = 0, /* Schur + direct Cholesky of the reduced system: in registers up to 160 unknowns,
                                        tile-sparse level-scheduled DMMA Cholesky above (fails if the factor needs > 24 GB) */
       BSNLS_LINEAR_ITERATIVE_SCHUR = 1, /* Schur + block-sparse PCG (block-Jacobi); in a sharded solve the block rows of the
                                            reduced system are partitioned over the ranks (the same rule as bsnls_shard_range) */
       BSNLS_LINEAR_AUTO = 2 };          /* direct; block-sparse PCG only if the tile factor would not fit */
typedef struct bsnls_options {
  int32_t max_num_iterations;            /* 50 */
  double max_solver_time_in_seconds;     /* 1e9 */
  double function_tolerance;             /* 1e-6 */
  double gradient_tolerance;             /* 1e-10 */
  double parameter_tolerance;            /* 1e-8 */
  double initial_trust_region_radius;    /* 1e4 */
  double max_trust_region_radius;        /* 1e16 */
  double min_trust_region_radius;        /* 1e-32 */
  double min_relative_decrease;          /* 1e-3 */
  double min_lm_diagonal;                /* 1e-6 */
  double max_lm_diagonal;                /* 1e32 */
  int32_t max_num_consecutive_invalid_steps; /* 5 */
  int32_t jacobi_scaling;                /* 1 */
  int32_t linear_solver;                 /* BSNLS_LINEAR_AUTO */
  int32_t pcg_max_iterations;            /* 500 */
  double pcg_tolerance;                  /* 1e-13: ||r|| <= tol*||b|| */
  int32_t device;                        /* CUDA ordinal, -1 = current */
} bsnls_options;

/* ceres::TerminationType values the caller's IsSolutionUsable() distinguishes
   (fixed_lag_smoother.cpp:286). */
enum { BSNLS_CONVERGENCE = 0, BSNLS_NO_CONVERGENCE = 1, BSNLS_FAILURE = 2 };

/* ceres::IterationSummary fields */
typedef struct bsnls_iteration {
  int32_t iteration;
  int32_t step_is_valid;
  int32_t step_is_successful;
  int32_t linear_solver_iterations;
  double cost;
  double cost_change;
  double gradient_max_norm;
  double gradient_norm;
  double step_norm;
  double relative_decrease;
  double trust_region_radius;
} bsnls_iteration;

/* ceres::Solver::Summary fields read by fixed_lag_smoother.cpp:286-292,693-717 */
typedef struct bsnls_summary {
  int32_t termination_type;
  int32_t num_iterations;         /* = iterations.size() */
  int32_t num_successful_steps;
  int32_t num_unsuccessful_steps;
  double initial_cost;
  double final_cost;
  double fixed_cost;
  double total_time_in_seconds;   /* host wall clock around the call, incl. H2D/D2H */
  double device_time_in_seconds;  /* CUDA events around the LM loop only */
  int32_t num_parameter_blocks_reduced, num_parameters_reduced, num_effective_parameters_reduced;
  int32_t num_residual_blocks_reduced, num_residuals_reduced;
  int32_t num_e_blocks, num_f_blocks, reduced_system_dim, linear_solver_used;
  int64_t reduced_system_nnz_blocks; /* 3x3 blocks of S in the lower triangle incl. diagonal */
  char message[160];
} bsnls_summary;

typedef struct bsnls_handle bsnls_handle;

void bsnls_default_options(bsnls_options* options);
/* replaces: ceres::Problem problem(problem_options_) (hash_graph.cpp:419) — one handle per graph */
int bsnls_create(const bsnls_options* options, bsnls_handle** handle);
void bsnls_destroy(bsnls_handle* handle);
const char* bsnls_last_error(const bsnls_handle* handle); /* NULL handle: last create error */

/* replaces: HashGraph::createProblem (hash_graph.cpp:475-522). Copies the flattened graph to the device,
   builds the elimination structure (landmark-sorted observations, reduced-system pattern). */
int bsnls_set_problem(bsnls_handle* handle, const bsnls_problem* problem);
/* re-upload parameter values only (same structure), e.g. between repeated solves of one window */
int bsnls_set_values(bsnls_handle* handle, const double* block_values);
int bsnls_set_options(bsnls_handle* handle, const bsnls_options* options);

/* replaces: ceres::Solve(options, &problem, &summary) (hash_graph.cpp:423, :443).
   iterations may be NULL; at most max_iterations_out records are written. On BSNLS_FAILURE the
   parameter values held by the handle are left as they were before the call. */
int bsnls_solve(bsnls_handle* handle, bsnls_summary* summary, bsnls_iteration* iterations,
                int32_t max_iterations_out);
/* replaces: the in-place update of Variable::data() by Ceres (hash_graph.cpp:421-423) */
int bsnls_get_values(bsnls_handle* handle, double* block_values);

/* replaces: ceres::Problem::Evaluate (hash_graph.cpp:446-455): cost, residuals (createProblem order, after
   the loss correction), gradient (tangent space, non-constant blocks in createProblem order). Any output
   pointer may be NULL. num_residuals / num_gradient report the lengths. */
int bsnls_evaluate(bsnls_handle* handle, double* cost, double* residuals, int64_t* num_residuals,
                   double* gradient, int64_t* num_gradient);

/* replaces: HashGraph::getCovariance (hash_graph.cpp:303-414) -> ceres::Covariance::Compute + GetCovarianceBlock[InTangentSpace]
   at the current values. pairs: num_pairs x 2 block indices; out: 16 doubles reserved per pair, the block row-major at the
   start of its slot — tangent space (fuse's default): 3 x 3; otherwise ambient size_i x size_j (quaternion rows / columns
   through the plus-Jacobian, as Ceres lifts them). Supported: pairs of non-constant blocks of the reduced system (poses,
   velocities, biases, extrinsics); a request for an eliminated landmark or a constant block fails with
   BSNLS_ERR_UNSUPPORTED, a rank-deficient problem with BSNLS_ERR_STATE (Ceres: Compute returns false). Direct solver only. */
int bsnls_get_covariance(bsnls_handle* handle, const int32_t* pairs, int32_t num_pairs, int32_t use_tangent_space, double* out);

/* ---- introspection used by the parity tests ("ordering and sparsity pattern bit-exact") ---- */
/* replaces: fuse_constraints::detail::linearize (fuse_constraints/src/marginalize_variables.cpp:281-399) for `count`
   residual blocks given by their createProblem numbers (res_index), at the values the handle holds: the loss-corrected
   residual (BSNLS_MARGINAL_MAX_ROWS doubles per block, unused rows zero) and the tangent-space Jacobian
   (BSNLS_MARGINAL_MAX_ROWS x 30 row-major per block: three columns per parameter block in the factor's block order, unused
   ones zero). One evaluation of residuals and Jacobians on the device serves the whole batch; marginalizeVariables of the
   host mirror (host/marginalize.cpp) is its caller. Single GPU. */
int bsnls_linearize(bsnls_handle* handle, int32_t count, const int32_t* residual_blocks, double* residuals, double* jacobians);

/* e-block / f-block lists in elimination order: indices into the caller's block arrays. */
int bsnls_get_ordering(bsnls_handle* handle, int32_t* e_blocks, int32_t* num_e, int32_t* f_blocks,
                       int32_t* num_f);
/* Lower-triangular (row >= col) 3x3-block pattern of the reduced system, as (row f-block position,
   col f-block position) pairs sorted by (row, col). Pass NULL pointers to query the count. */
int bsnls_get_reduced_pattern(bsnls_handle* handle, int32_t* rows, int32_t* cols, int64_t* count);
/* Linearise at the current values with the given trust-region radius and return the dense reduced system
   (full symmetric matrix, column-major, dim x dim, in the program order of bsnls_get_ordering) and its right-hand
   side. For tests: the matrix is expanded on the host whatever the internal storage is. */
int bsnls_get_reduced_system(bsnls_handle* handle, double radius, double* S, double* rhs, int32_t dim);

/* Same system on its structural pattern, for graphs whose dense matrix does not fit: the lower-triangular 3x3 blocks in the
   order of bsnls_get_reduced_pattern (rows / cols are filled as that call fills them), 9 doubles per block, row-major.
   Pass NULL pointers to query the count. */
int bsnls_get_reduced_blocks(bsnls_handle* handle, double radius, int32_t* rows, int32_t* cols, double* vals,
                             int64_t* count, double* rhs);
/* One Levenberg-Marquardt step from a given point, for teacher-forced parity checks (no counterpart in the reference:
   ceres::Solve does not expose single steps). Uploads `values` (NULL: keep the current ones), starts from
   trust-region radius `radius`, evaluates (record 0: cost and gradient norms at the point), takes one step (record 1:
   step norm, cost change, relative decrease, the radius after the update). keep_jacobi_scale != 0 reuses the column
   scaling of the previous solve / trial step on this handle (Ceres freezes it at iteration 0). num_out < 2 when the
   step ended the minimisation (tolerance reached). An accepted step stays applied to the handle's values. */
int bsnls_trial_step(bsnls_handle* handle, const double* values, double radius, int32_t keep_jacobi_scale,
                     bsnls_iteration* out, int32_t* num_out);

/* ---- per-kernel entry points (one pass of the hot path; what bench.py's roofline times) ---- */
/* K1-K4: evaluate every factor (residual, tangent Jacobian, loss correction), cost, gradient. */
int bsnls_run_eval(bsnls_handle* handle, double* cost);
/* K5-K6: landmark elimination + reduced-system assembly for the given radius (needs a prior eval). */
int bsnls_run_schur(bsnls_handle* handle, double radius);
/* One complete linearisation as an LM iteration performs it: K1-K6 in one go (on large shards the chunk-major fused
   kernel evaluates the Jacobians, eliminates the landmarks and accumulates the reduced system without writing the
   intermediate records). cost (optional) = cost at the current values. bsnls_last_kernel_ms then reports the whole pass as
   eval_ms (schur_ms = 0). */
int bsnls_run_pass(bsnls_handle* handle, double radius, double* cost);
/* on != 0: bsnls_run_pass launches its kernels one by one with CUDA events between them, so that
   bsnls_last_breakdown_ms reports the split; otherwise (default) the pass is one CUDA-graph launch and only the total is
   timed. */
int bsnls_enable_breakdown(bsnls_handle* handle, int32_t on);
/* milliseconds spent in the last bsnls_run_eval / bsnls_run_schur, by CUDA events on the handle's stream */
int bsnls_last_kernel_ms(bsnls_handle* handle, float* eval_ms, float* schur_ms);
/* per-kernel split of the last run_eval + run_schur, milliseconds:
   [0] reprojection eval, [1] generic-factor eval, [2] landmark elimination, [3] gradient / column norms,
   [4] reduced-system clear, [5] pair (S assembly), [6] generic-factor S rows + damping,
   [7] NCCL exchange of the reduced system (sharded solve; 0 on one GPU) */
int bsnls_last_breakdown_ms(bsnls_handle* handle, float out[8]);
/* number of kernel launches issued by this handle since creation */
int64_t bsnls_launch_count(const bsnls_handle* handle);

/* ---- allocation-free entry point ----
   No counterpart in the reference (Ceres allocates freely on the host). It exists for callers that may not
   request device memory: real-time threads that cannot afford the device-wide synchronisation of
   cudaMalloc/cudaFree, processes whose GPU memory is a fixed arena, code inside a stream capture.
   Computes sum_i 0.5*rho(|r_i|^2) over `num_observations` reprojection observations (the solver's candidate-cost
   evaluation). All pointers are DEVICE pointers owned by the caller: d_values = concatenated parameter blocks;
   d_q_offset / d_t_offset / d_p_offset = per observation, offset into d_values of the orientation (4), position
   (3) and landmark (3) block; d_camera_index into d_cameras (BSNLS_CAMERA_STRIDE doubles each); d_pixel 2 per
   observation; d_cost one double. `stream` is a cudaStream_t (NULL = default stream); the call only enqueues.
   Scratch is a fixed 128 KiB array inside the library, shared by all calls on a device (calls that overlap on the
   GPU race), resident on every GPU that loads the library; its address is looked up on the first call per
   device, which should precede any stream capture. The device is the CURRENT one (cudaGetDevice) at the time of the
   call: make the device that owns `stream` and the buffers current before calling. More than BSNLS_NOALLOC_MAX_OBSERVATIONS observations are
   refused (BSNLS_ERR_INVALID); error_message (optional, error_capacity bytes) receives the reason. */
enum { BSNLS_NOALLOC_MAX_OBSERVATIONS = 4194304 };
int bsnls_reprojection_cost_noalloc(const double* d_values, const int32_t* d_q_offset, const int32_t* d_t_offset,
                                    const int32_t* d_p_offset, const int32_t* d_camera_index, const double* d_pixel,
                                    const double* d_cameras, int32_t num_observations, double* d_cost, void* stream,
                                    char* error_message, int32_t error_capacity);

/* ---- snapshot: the flattened problem on disk (host only, no device needed) ----
   No replayable counterpart in the reference: its Boost serialisation keeps only pixel_ of the visual constraints
   (bs_constraints/include/bs_constraints/visual/euclidean_reprojection_constraint.h:99-103), and constraint UUIDs are
   random per run, so a graph read back from a bag cannot be solved again in the same order. The snapshot stores the SoA
   buffers of bsnls_problem as the flattener produced them (createProblem order), so that the GPU path, the CPU checker
   and a Ceres adapter consume identical arrays. Format: csrc/snapshot.cpp. bsnls_problem_load allocates; release with
   bsnls_problem_free. */
int bsnls_problem_save(const bsnls_problem* problem, const char* path);
int bsnls_problem_load(const char* path, bsnls_problem** problem);
void bsnls_problem_free(bsnls_problem* problem);

/* ---- multi-GPU: landmark-sharded solve, one process per GPU ---- */
/* Balanced cut of an ordered landmark list: rank r gets the entries [lo, hi), chosen so that every rank holds about the
   same number of observations (lm_ptr: prefix sums of observations per landmark in that order, num_landmarks + 1
   entries). Host-only helper, no device needed. The sharded solve applies it to the landmarks in LOCALITY order (by the
   earliest pose along the trajectory that observes them), so that a rank works on one stretch of the trajectory. */
int bsnls_shard_range(const int32_t* lm_ptr, int32_t num_landmarks, int32_t rank, int32_t world_size, int32_t* lo, int32_t* hi);
/* The landmarks this rank keeps after bsnls_set_problem: positions in the elimination order of bsnls_get_ordering
   (all of them on a single GPU). Pass NULL to query the count. The shards of all ranks partition the e-blocks. */
int bsnls_get_shard(bsnls_handle* handle, int32_t* e_positions, int32_t* count);
/* 128-byte NCCL unique id, generated on rank 0 and distributed by the caller (torch.distributed). */
int bsnls_nccl_unique_id(char id[128]);
/* Collective over all ranks. Every rank then calls bsnls_set_problem with the SAME full problem; the
   library keeps only the landmarks (and their observations) of its shard. */
int bsnls_create_sharded(const bsnls_options* options, int32_t rank, int32_t world_size,
                         const char nccl_id[128], bsnls_handle** handle);

#ifdef __cplusplus
}
#endif
#endif /* BSNLS_H_ */
