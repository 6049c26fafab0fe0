/*
 * mini_opt_b200 — C ABI of the batched B200 solver for mini_opt's hot path.
 *
 * The reference (gareth-cross/mini_opt) has no FFI: its boundary is the C++ class surface
 *   QPInteriorPointSolver::Setup/Solve            include/mini_opt/qp.hpp:167-205, source/qp.cc:20-151
 *   QPInteriorPointSolver::Iterate (private)      source/qp.cc:153-201
 *   ConstrainedNonlinearLeastSquares::LinearizeAndFillQP / Solve   source/nonlinear.cc:75-214
 * Every entry point below names the reference function it replaces. Conventions:
 *   - plain C structs, caller-owned buffers, no exceptions; `int` return (MO_OK or a negative mo_error);
 *   - per-problem numerical trouble never aborts a batch: it is reported in `status[i]`;
 *   - unless the name ends in `_host`, every data pointer is a DEVICE pointer valid on `device`,
 *     and the call is asynchronous on `cuda_stream` (a `cudaStream_t`, NULL = legacy default stream);
 *   - all floating point data is IEEE fp64; matrices are column-major like Eigen::MatrixXd.
 */
#ifndef MINI_OPT_B200_H_
#define MINI_OPT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MO_ABI_VERSION 2

/* ---- return codes of the entry points (not per-problem status) ---- */
typedef enum mo_error {
  MO_OK = 0,
  MO_ERR_INVALID_ARGUMENT = -1, /* NULL where data is required, negative dims, constraint index out of range */
  MO_ERR_INVALID_PARAMS = -2,   /* the reference's CheckParams would have thrown (qp.cc:76-82, nonlinear.cc:49-73) */
  MO_ERR_UNSUPPORTED_SHAPE = -3,/* problem too large for every kernel variant (shared memory) */
  MO_ERR_CUDA = -4,             /* a CUDA runtime call failed; see mo_last_error_string() */
  MO_ERR_NO_DEVICE = -5         /* no CUDA device / wrong architecture (needs sm_100) */
} mo_error;

/* ---- per-problem termination status of the interior point solve ----
 * 0/1 are QPInteriorPointTerminationState (include/mini_opt/structs.hpp:97-102).
 * 2 replaces the FailedFactorization exception (source/qp.cc:303-307) and
 * 3 replaces F_ASSERT(all s > 0) (source/qp.cc:285): a batch cannot throw per problem. */
enum {
  MO_QP_SATISFIED_KKT_TOL = 0,
  MO_QP_MAX_ITERATIONS = 1,
  MO_QP_FAILED_FACTORIZATION = 2,
  MO_QP_NONPOSITIVE_SLACK = 3
};

/* BarrierStrategy, include/mini_opt/structs.hpp:24-31 */
enum { MO_BARRIER_COMPLEMENTARITY = 0, MO_BARRIER_FIXED_DECREASE = 1, MO_BARRIER_PREDICTOR_CORRECTOR = 2 };
/* InitialGuessMethod, include/mini_opt/structs.hpp:34-41 */
enum { MO_GUESS_NAIVE = 0, MO_GUESS_SOLVE_EQUALITY_CONSTRAINED = 1, MO_GUESS_USER_PROVIDED = 2 };

/* flags of mo_qp_batch */
enum {
  /* Caller promises m == 2n and constraint 2i is (i,+1,b), constraint 2i+1 is (i,-1,b), i.e.
   * Var(i) >= lo_i, Var(i) <= hi_i with ineq_b = {-lo_0, hi_0, -lo_1, hi_1, ...}. ineq_var / ineq_a
   * are then not read (may be NULL). This is the box layout of BASELINE configs 2-5. */
  MO_QP_FLAG_BOX_PAIRS = 1,
  /* HOST entry points only (mo_qp_ip_solve_batch_host / _host_multi): compact input formats that cut the bytes
   * crossing PCIe; they are expanded on the device into the layout above before the solve.
   * PACKED_G: G is [batch][n(n+1)/2], the lower triangle column by column (column j holds rows j..n-1) -- exactly the
   * entries the reference reads (qp.cc:289, :404); 288 instead of 512 bytes per problem at n = 8.
   * UNIFORM_BOX (needs BOX_PAIRS): ineq_b is ONE [2n] vector {-lo_0, hi_0, ...} shared by the whole batch. */
  MO_QP_FLAG_PACKED_G = 2,
  MO_QP_FLAG_UNIFORM_BOX = 4
};

/*
 * A batch of QPs of uniform shape (struct QP, include/mini_opt/qp.hpp:104-124):
 *    minimise 1/2 x'Gx + c'x   s.t.  A_eq x + b_eq = 0,   a_i * x[var_i] + b_i >= 0.
 * Problem p occupies [p*stride, (p+1)*stride) of each array.
 */
typedef struct mo_qp_batch {
  int32_t n;              /* variables per problem (N)               */
  int32_t k;              /* equality rows per problem (K), may be 0 */
  int32_t m;              /* inequality slots per problem (M), may be 0 */
  int32_t flags;          /* MO_QP_FLAG_*                            */
  int64_t batch;          /* number of problems                      */
  const double* G;        /* [batch][n*n] column-major; only the lower triangle is read (qp.cc:289,404) */
  const double* c;        /* [batch][n]                              */
  const double* A_eq;     /* [batch][k*n] column-major k x n (NULL if k == 0) */
  const double* b_eq;     /* [batch][k]                              */
  const int32_t* ineq_var;/* [batch][m] LinearInequalityConstraint::variable (qp.hpp:28-35) */
  const double* ineq_a;   /* [batch][m]                              */
  const double* ineq_b;   /* [batch][m]                              */
  const int32_t* ineq_count; /* optional [batch]: problem p uses only its first ineq_count[p] <= m slots
                                (ragged batches such as GenerateRandomQP, test/qp_test.cc:483-524). NULL = m. */
} mo_qp_batch;

/* QPInteriorPointSolver::Params, include/mini_opt/qp.hpp:134-164 (same defaults via mo_qp_params_default). */
typedef struct mo_qp_params {
  double initial_mu;
  double sigma;
  double termination_kkt_tol;
  double termination_complementarity_tol;
  int32_t max_iterations;
  int32_t barrier_strategy;
  int32_t decrease_mu_only_on_small_error;
  int32_t initial_guess_method;
  int32_t initialize_mu_with_complementarity;
  int32_t reserved;
} mo_qp_params;

#define MO_QP_ITER_LOG_STRIDE 14
/* One row of iter_log = QPInteriorPointIteration (include/mini_opt/structs.hpp:81-94):
 *  [0..3]  kkt_initial {r_dual, r_comp, r_primal_eq, r_primal_ineq}
 *  [4..7]  kkt_final   {same}
 *  [8]     ip_outputs.mu   [9] alpha.primal [10] alpha.dual
 *  [11]    alpha_probe.primal [12] alpha_probe.dual [13] mu_affine   (NaN unless PREDICTOR_CORRECTOR) */

typedef struct mo_qp_out {
  double* vars;      /* [batch][n+2m+k] = [x | s | y | z] (qp.cc:548-582). In/out for MO_GUESS_USER_PROVIDED.
                        HOST entry points: may be NULL when `x` is given (only x_block() is copied back). */
  int32_t* status;   /* [batch] MO_QP_*                                                  */
  int32_t* iters;    /* [batch] == outputs.iterations.size()                             */
  double* kkt_final; /* optional [batch][4]: kkt_final of the last iteration              */
  double* lagrange;  /* optional [batch][2]: {min y, ||y||_inf} (qp.cc:539-546); NaN if k == 0 */
  double* iter_log;  /* optional [batch][max_iterations][MO_QP_ITER_LOG_STRIDE]           */
  double* x;         /* optional [batch][n]: x_block() alone (qp.hpp:187). HOST entry points only: with vars == NULL
                        the device-to-host traffic per problem drops from 8(n+2m+k) to 8n bytes.     */
} mo_qp_out;

/* Fills *p with the reference defaults (qp.hpp:136-163). */
void mo_qp_params_default(mo_qp_params* p);

/* Elements of `vars` per problem: n + 2m + k. */
int64_t mo_qp_vars_stride(int32_t n, int32_t k, int32_t m);

/*
 * Replaces: QPInteriorPointSolver::Setup + Solve for every problem of the batch
 * (source/qp.cc:20-73, :100-151; ComputeInitialGuess :439-482).
 * Device pointers, asynchronous on `cuda_stream`.
 */
int mo_qp_ip_solve_batch(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out,
                         int device, void* cuda_stream);

/*
 * Same contract with HOST pointers: stages through pinned memory, H2D, solve, D2H, synchronises
 * before returning. This is the call a host-side user of the reference makes (data lives in Eigen
 * objects on the host) and what bench.py's `e2e` times.
 */
int mo_qp_ip_solve_batch_host(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out,
                              int device);

/*
 * Same as mo_qp_ip_solve_batch_host over SEVERAL devices of one node from one process: the batch is split into
 * n_devices contiguous shards, one host thread and its own stream ring per device, no collective, every shard's results
 * written straight into the caller's single output arrays ("gathered only at the end": the gather is the D2H copies).
 * `devices[i]` are CUDA device ordinals. Returns the first error of any shard.
 */
int mo_qp_ip_solve_batch_host_multi(const mo_qp_batch* problems, const mo_qp_params* params, const mo_qp_out* out,
                                    const int* devices, int n_devices);

/* Pinned (page-locked) host memory: buffers allocated here are DMA'd directly by the *_host entry points;
 * pageable buffers work too but are staged by the driver. */
void* mo_host_alloc(size_t bytes);
/* The same, with the pages interleaved over all NUMA nodes of the host: for buffers that feed several GPUs from one process
 * (mo_qp_ip_solve_batch_host_multi). Falls back to mo_host_alloc where the policy cannot be set. */
void* mo_host_alloc_interleaved(size_t bytes);
void mo_host_free(void* p);

/*
 * Replaces: one call of QPInteriorPointSolver::Iterate(mu, strategy) (source/qp.cc:153-201) on
 * caller-provided state `vars` [batch][n+2m+k] (updated in place). Used for step-level parity with
 * the reference's Elimination tests (test/qp_test.cc:101-241).
 *   delta_out     optional [batch][n+2m+k]: the Newton direction [dx|ds|dy|dz] before scaling by alpha
 *   ip_out        optional [batch][6]: {mu, alpha.primal, alpha.dual, alpha_probe.primal, alpha_probe.dual, mu_affine}
 *   kkt_out       optional [batch][4]: KKT error norms (ComputeErrors(mu), qp.cc:423-437) AFTER the step
 *   status        [batch]: MO_QP_SATISFIED_KKT_TOL (0) if the step was taken, MO_QP_FAILED_FACTORIZATION /
 *                 MO_QP_NONPOSITIVE_SLACK otherwise (state untouched)
 */
int mo_qp_ip_newton_step_batch(const mo_qp_batch* problems, double mu, int32_t barrier_strategy,
                               double* vars, double* delta_out, double* ip_out, double* kkt_out,
                               int32_t* status, int device, void* cuda_stream);

/* QPNullSpaceTerminationState, include/mini_opt/structs.hpp:137-142; 2 = rank(A_eq) < k (batched stand-in: the reference
 * multiplies mismatched blocks there). */
enum { MO_QP_NULLSPACE_SUCCESS = 0, MO_QP_NULLSPACE_NOT_POSITIVE_DEFINITE = 1, MO_QP_NULLSPACE_RANK_DEFICIENT = 2 };

/*
 * Replaces: QPNullSpaceSolver::Solve (source/qp.cc:679-729) for every problem of the batch: equality-constrained QP by
 * column-pivoted Householder QR of A_eq^T, reduced Hessian Q2' G Q2, Cholesky. Reads G, c, A_eq, b_eq (k > 0); the
 * inequality fields are ignored. x [batch][n] (zeros unless SUCCESS), status [batch] MO_QP_NULLSPACE_*. Device
 * pointers, asynchronous on cuda_stream. This is the solver the NLS driver selects for problems with equality
 * constraints and no inequalities (source/nonlinear.cc:83-89); mo_nls_solve_chain_ik_batch does the same.
 */
int mo_qp_nullspace_solve_batch(const mo_qp_batch* problems, double* x, int32_t* status, int device, void* cuda_stream);
/* Same with HOST pointers (H2D, solve, D2H; synchronous): the call QPNullSpaceSolver of the C++ mirror makes. */
int mo_qp_nullspace_solve_batch_host(const mo_qp_batch* problems, double* x, int32_t* status, int device);

/*
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation). Same contract and results as
 * mo_qp_nullspace_solve_batch, but the per-warp workspace (QR factors, Q, reduced Hessian: O(n^2) doubles per resident
 * problem) lives in ONE file-scope device array of 32 MiB instead of dynamic shared memory, and the call makes no
 * cudaMalloc / cudaFree / allocating library call. Limits and caveats:
 *   - n <= MO_NOALLOC_MAX_N (else MO_ERR_UNSUPPORTED_SHAPE with a message in mo_last_error_string()); any batch size;
 *   - every GPU that loads the library holds the 32 MiB array until the process exits;
 *   - every call on a device shares the array: calls that OVERLAP on the GPU (different streams, no ordering) race;
 *     order them on one stream or with events. Nothing in it needs to start at zero (every word is written before it is
 *     read), so a call does not clear it and it keeps the last call's contents.
 */
#define MO_NOALLOC_MAX_N 128
int mo_qp_nullspace_solve_batch_noalloc(const mo_qp_batch* problems, double* x, int32_t* status, int device,
                                        void* cuda_stream);

/* =====================================================================================================
 * Nonlinear least squares: Gauss-Newton / Levenberg-Marquardt linearization and the SQP loop
 * (ConstrainedNonlinearLeastSquares, include/mini_opt/nonlinear.hpp, source/nonlinear.cc)
 * ===================================================================================================== */

/* LineSearchStrategy, include/mini_opt/structs.hpp:146-151 */
enum { MO_LS_ARMIJO_BACKTRACK = 0, MO_LS_POLYNOMIAL_APPROXIMATION = 1 };
/* Inner QP solver of mo_nls_solve_chain_ik_batch: AUTO = warp-synchronous register solver (unpivoted LDL') when
 * n + k <= 16, m <= 32 and no variable carries more than two constraints, else GENERIC; GENERIC = the Eigen-pivot-order
 * solver of the generic QP kernel (audit path: reproduces the reference's QP iteration counts most closely). */
enum { MO_NLS_QP_AUTO = 0, MO_NLS_QP_GENERIC = 1 };
/* NLSTerminationState, include/mini_opt/structs.hpp:233-248 */
enum {
  MO_NLS_MAX_ITERATIONS = 0,
  MO_NLS_SATISFIED_ABSOLUTE_TOL = 1,
  MO_NLS_SATISFIED_RELATIVE_TOL = 2,
  MO_NLS_SATISFIED_FIRST_ORDER_TOL = 3,
  MO_NLS_MAX_LAMBDA = 4,
  MO_NLS_QP_INDEFINITE = 5,
  MO_NLS_USER_CALLBACK = 6,
  /* batched stand-ins for exceptions: the inner QP threw FailedFactorization / hit F_ASSERT(s > 0) */
  MO_NLS_QP_FAILED = 7
};
/* Retraction (nonlinear.hpp:127). 0: x + alpha dx (nonlinear.cc:166).  1: ModPi(x + alpha dx)
 * (geometry_utils/angle_utils.hpp:9-17), the retraction every actuator-chain test installs
 * (test/nonlinear_test.cc:864-870, 1076-1082). */
enum { MO_RETRACT_ADD = 0, MO_RETRACT_MOD_PI = 1 };

/* ConstrainedNonlinearLeastSquares::Params, include/mini_opt/nonlinear.hpp:37-124 (same defaults via
 * mo_nls_params_default). log_qp_eigenvalues has no batched equivalent. */
typedef struct mo_nls_params {
  int32_t max_iterations;
  int32_t max_qp_iterations;
  double termination_kkt_tolerance;
  double absolute_exit_tol;
  double relative_exit_tol;
  double absolute_first_derivative_tol;
  int32_t max_line_search_iterations;
  int32_t line_search_strategy;
  double armijo_search_tau;
  double equality_penalty_initial;
  double equality_penalty_scale_factor;
  double equality_penalty_rho;
  double lambda_initial;
  double lambda_failure_init;
  double lambda_decrease_on_success;
  double lambda_decrease_on_restore;
  double max_lambda;
  double min_lambda;
  int32_t retraction; /* MO_RETRACT_* */
  int32_t inner_qp_solver; /* MO_NLS_QP_* */
} mo_nls_params;

void mo_nls_params_default(mo_nls_params* p);

/*
 * Replaces: ConstrainedNonlinearLeastSquares::LinearizeAndFillQP (source/nonlinear.cc:170-214) with
 * Residual::Model::UpdateHessian / UpdateJacobian (include/mini_opt/residual.hpp:186-250) for problems whose
 * residuals are supplied as dense per-problem Jacobians evaluated at x:
 *   G  = sum over cost residual blocks of J_b' J_b  (lower triangle; strict upper triangle exactly 0)  + lambda I
 *   c  = sum J_b' r_b,   f = sum 1/2 |r_b|^2,   A_eq = J_eq, b_eq = r_eq, equality = |r_eq|_1,
 *   constraint i shifted to the linearization point: (v, a, a x_v + b)   (qp.hpp:57-59)
 * All pointers are DEVICE pointers; asynchronous on cuda_stream.
 */
typedef struct mo_nls_linearize_args {
  int32_t n;          /* variables                                                      */
  int32_t r_cost;     /* total cost rows                                                */
  int32_t cost_block; /* rows per cost Residual (r_cost % cost_block == 0; 0 -> one block) */
  int32_t r_eq;       /* equality rows (k of the QP)                                    */
  int32_t m;          /* inequality constraints                                         */
  int32_t reserved;
  int64_t batch;
  const double* x;        /* [batch][n]                                  */
  const double* J_cost;   /* [batch][r_cost*n] column-major r_cost x n   */
  const double* r_cost_v; /* [batch][r_cost]                             */
  const double* J_eq;     /* [batch][r_eq*n] column-major (NULL if r_eq == 0) */
  const double* r_eq_v;   /* [batch][r_eq]                               */
  const double* lambda;   /* [batch] LM damping per problem              */
  const int32_t* ineq_var;/* [batch][m]                                  */
  const double* ineq_a;   /* [batch][m]                                  */
  const double* ineq_b;   /* [batch][m]                                  */
  /* outputs: exactly the arrays of a mo_qp_batch with (n, k = r_eq, m) */
  double* G;              /* [batch][n*n]                                */
  double* c;              /* [batch][n]                                  */
  double* A_eq;           /* [batch][r_eq*n]                             */
  double* b_eq;           /* [batch][r_eq]                               */
  double* ineq_b_shifted; /* [batch][m]                                  */
  double* errors;         /* [batch][2] Errors{f, equality} (structs.hpp:169-186) */
} mo_nls_linearize_args;

int mo_nls_linearize_batch(const mo_nls_linearize_args* args, int device, void* cuda_stream);

/*
 * The transform-chain residual family (test/transform_chains.hpp:118-211): every problem of the batch is one
 * ActuatorChain of `links` links with the same activity masks. Link i is Pose(q_i, t_i) (parent_T_child); the
 * active Euler angles (XYZ composition, so3.hpp:293-338) and translations of all links, in link order, are the
 * n optimisation variables. Residual rows are linear functionals of the effector translation:
 *     row = w . root_t_effector - target
 * (the y_res / x_eq residuals of test/nonlinear_test.cc:840-859). Inequalities are (var, a, b) triplets shared
 * by the whole batch (joint limits).
 */
typedef struct mo_chain_ik_batch {
  int32_t links;
  int32_t n;          /* = number of non-zero mask entries                                   */
  int32_t n_cost;     /* cost rows                                                            */
  int32_t cost_block; /* rows per cost Residual (0 -> one block)                              */
  int32_t n_eq;       /* equality rows                                                        */
  int32_t m;          /* inequality constraints                                               */
  int64_t batch;
  const double* link_pose;   /* [links][7] = (qw, qx, qy, qz, tx, ty, tz), shared by the batch  */
  const uint8_t* link_mask;  /* [links][6] = active {rx, ry, rz, tx, ty, tz}                     */
  const double* cost_w;      /* [n_cost][3], shared                                              */
  const double* cost_target; /* [batch][n_cost]                                                  */
  const double* eq_w;        /* [n_eq][3], shared (NULL if n_eq == 0)                            */
  const double* eq_target;   /* [batch][n_eq]                                                    */
  const int32_t* ineq_var;   /* [m] shared                                                       */
  const double* ineq_a;      /* [m]                                                              */
  const double* ineq_b;      /* [m]                                                              */
  const double* x0;          /* [batch][n] initial variables                                     */
} mo_chain_ik_batch;

typedef struct mo_nls_out {
  double* x;              /* [batch][n] nls.variables() after Solve                                  */
  int32_t* termination;   /* [batch] MO_NLS_*                                                        */
  int32_t* iterations;    /* [batch] outputs.iterations.size()                                       */
  int32_t* qp_iterations; /* [batch] total interior point iterations (Newton steps) over all QPs     */
  int32_t* ls_steps;      /* [batch] total line search evaluations (NUM_LINE_SEARCH_STEPS)           */
  double* errors;         /* optional [batch][2]: Errors{f, equality} at the returned x              */
} mo_nls_out;

/*
 * Replaces: ActuatorChain::Update (test/transform_chains.cc:165-244) for `batch` parameter vectors.
 *   params [batch][n];  translation [batch][3];  rotation_q [batch][4] (w,x,y,z);  t_D_params, r_D_params
 *   [batch][3*n] column-major 3 x n. Any output may be NULL. The chain description comes from `chain`
 *   (links, n, batch, link_pose, link_mask; the other fields are ignored).
 */
int mo_chain_fk_batch(const mo_chain_ik_batch* chain, const double* params, double* translation,
                      double* rotation_q, double* t_D_params, double* r_D_params, int device, void* cuda_stream);

/*
 * Replaces: ConstrainedNonlinearLeastSquares::Solve (source/nonlinear.cc:75-158) -- LinearizeAndFillQP,
 * ComputeStepDirection (interior point QP, :216-259), ComputeQPCostDerivative, SelectPenalty, SelectStepSize,
 * UpdateLambdaAndCheckExitConditions -- for every problem of a transform-chain batch, entirely on the device.
 * Problems with equality constraints and NO inequalities use the null-space solver like the reference
 * (nonlinear.cc:83-89), with the penalty update of inequality (18.36) (nonlinear.cc:492-500) and QP_INDEFINITE when the
 * reduced Hessian is not positive definite. Device pointers, asynchronous on cuda_stream.
 */
int mo_nls_solve_chain_ik_batch(const mo_chain_ik_batch* problems, const mo_nls_params* params,
                                const mo_nls_out* out, int device, void* cuda_stream);

/*
 * Device-side residual FAMILIES defined by functors: the reference's MakeResidual<R, P>(index, functor)
 * (include/mini_opt/residual.hpp:121-143) keeps user functors on the host; here a problem class is a struct of two
 * __device__ functions (mini_opt_b200/csrc/nls_families.cuh) for which the SQP + LM kernel is instantiated, so the whole
 * ConstrainedNonlinearLeastSquares::Solve loop (source/nonlinear.cc:75-158) runs on the device for every start point of a
 * batch. Built in: the reference's own test problems (test/nonlinear_test.cc:375-387, 580-594, 721-760).
 */
enum { MO_NLS_FAMILY_ROSENBROCK = 0, MO_NLS_FAMILY_HIMMELBLAU = 1, MO_NLS_FAMILY_SPHERE_PRODUCT = 2 };

typedef struct mo_nls_family_batch {
  int32_t family;          /* MO_NLS_FAMILY_*                                                        */
  int32_t m;               /* inequality constraints, (var, a, b) triplets shared by the batch        */
  int64_t batch;
  const int32_t* ineq_var; /* [m]                                                                    */
  const double* ineq_a;    /* [m]                                                                    */
  const double* ineq_b;    /* [m]                                                                    */
  const double* x0;        /* [batch][n] start points                                                */
  const double* data;      /* optional [batch][data_stride]: per-problem constants of the functor     */
  int32_t data_stride;
  int32_t reserved;
} mo_nls_family_batch;

/* n, cost rows and equality rows of a family (any output may be NULL). */
int mo_nls_family_dims(int32_t family, int32_t* n, int32_t* n_cost, int32_t* n_eq);
/* Device pointers, asynchronous on cuda_stream. mo_nls_out.x is [batch][n of the family]. */
int mo_nls_solve_family_batch(const mo_nls_family_batch* problems, const mo_nls_params* params, const mo_nls_out* out,
                              int device, void* cuda_stream);

/* Name of the kernel variant mo_qp_ip_solve_batch would launch for this shape ("warp16", "cta", ...). */
const char* mo_qp_kernel_variant(int32_t n, int32_t k, int32_t m, int32_t flags);

/* Number of kernel launches issued by this library since load (bench.py's gpu_launches). */
int64_t mo_kernel_launch_count(void);

/* Human-readable description of the last MO_ERR_CUDA in this thread. */
const char* mo_last_error_string(void);

/* Measures this GPU's FP64 FMA peak (dependent-free DFMA streams, CUDA events, best of 5) in TFLOP/s: the
 * roofline denominator for the FP64-bound kernels (bench.py). Synchronous. */
int mo_measure_fp64_peak(int device, double* tflops_out);
/* Same for the FP64 tensor path (mma.sync.m8n8k4.f64 = SASS DMMA.8x8x4, 8 independent accumulators per warp): the
 * denominator for the kernels whose O(p^3) work runs on DMMA (n >= 16 factorizations, n >= 32 linearization). */
int mo_measure_dmma_peak(int device, double* tflops_out);

int mo_abi_version(void);

/* Measurement hook (profiles/, tests): staging mode of the kernels that load per-problem arrays with TMA bulk copies
 * (cp.async.bulk + mbarrier). 1 = TMA (default), 0 = plain per-lane loads. Returns the previous value. Process-wide. */
int mo_debug_set_tma(int enabled);
/* Measurement hook: refill of the thread-per-problem kernel through cp.async one round ahead (1, default) or through
 * registers (0, the round-1 schedule). Results are bit-identical. Returns the previous value. */
int mo_debug_set_tpp_async(int enabled);
/* Measurement / cross-check hook: 0 = default kernel for the DMMA shapes (warp per problem, "dmma_warp"), 1 = the
 * CTA-per-problem kernel ("dmma_cta"). Returns the previous value. Process-wide; the library reads no environment. */
int mo_debug_set_dm_kernel(int which);
/* Measurement hook: the warp-per-problem DMMA kernel in lock-step CTAs (1; measured slower, profiles/r02_wd_lockstep_ab.txt)
 * or one-warp CTAs (0, default). Bit-identical. */
int mo_debug_set_wd_lock(int enabled);
/* Measurement hooks of the lock-step chain-IK kernel: warps per CTA (1, 16, 24; -1 on anything else) and which of the
 * two optional CTA barriers of a pass are taken (bit 0: after the linearization, bit 1: after the inner QP; bits 2-3: the
 * opening barrier only every 2nd / 4th pass). Results are
 * bit-identical for every value. Each returns the previous value. */
int mo_debug_set_ik_warps(int warps);
int mo_debug_set_ik_sync(int mask);

#ifdef __cplusplus
} /* extern "C" */
#endif
#endif /* MINI_OPT_B200_H_ */
