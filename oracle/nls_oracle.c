/*
 * TEST INFRASTRUCTURE — not part of the product. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product path never does.
 *
 * Plain-C restatement of the reference's nonlinear least squares path (gareth-cross/mini_opt):
 *   Residual::Model::UpdateHessian / UpdateJacobian      include/mini_opt/residual.hpp:186-250
 *   LinearizeAndFillQP                                   source/nonlinear.cc:170-214
 *   ConstrainedNonlinearLeastSquares::Solve and helpers  source/nonlinear.cc:75-158, 216-605
 *   ActuatorChain::Update / ComputeChain / ActuatorLink  test/transform_chains.cc:23-258
 *   SO3FromEulerAngles_ (XYZ), EulerAnglesFromSO3        geometry_utils/so3.hpp:293-338, 365-371
 *   ModPi                                                geometry_utils/angle_utils.hpp:9-17
 * including the Eigen 3.4.0 quaternion operations those lean on (product, rotation of a vector,
 * toRotationMatrix). The inner QP is solved by the QP oracle of this directory (qp_ip_oracle.c).
 *
 * Parity: PINNED. tests/test_oracle_nls.py checks this file against the reference itself
 * (oracle/_ref/libminiopt_ref.so) and against the golden vectors under tests/golden/ recorded from it.
 */
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "mini_opt_b200.h"

int64_t oracle_qp_solve_batch(const mo_qp_batch* b, const mo_qp_params* prm, const mo_qp_out* out);
int oracle_qp_nullspace_one(int n, int k, const double* G, const double* c, const double* A, const double* b, double* x);

#define MAX_LINKS 64

/* ---------------------------------------------------------------------------------------------------
 * Eigen::Quaternion operations, element order (w, x, y, z)
 * --------------------------------------------------------------------------------------------------- */
typedef struct { double w, x, y, z; } quat;

static quat quat_mul(quat a, quat b) { /* Eigen/src/Geometry/Quaternion.h quat_product */
  quat r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  return r;
}

static quat quat_conj(quat a) { quat r = {a.w, -a.x, -a.y, -a.z}; return r; }

static void quat_rotate(quat q, const double* v, double* out) { /* QuaternionBase::_transformVector */
  double uv[3] = {q.y * v[2] - q.z * v[1], q.z * v[0] - q.x * v[2], q.x * v[1] - q.y * v[0]};
  uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
  const double c[3] = {q.y * uv[2] - q.z * uv[1], q.z * uv[0] - q.x * uv[2], q.x * uv[1] - q.y * uv[0]};
  out[0] = v[0] + q.w * uv[0] + c[0];
  out[1] = v[1] + q.w * uv[1] + c[1];
  out[2] = v[2] + q.w * uv[2] + c[2];
}

/* R column-major 3x3: R[c*3+r] (QuaternionBase::toRotationMatrix) */
static void quat_matrix(quat q, double* R) {
  const double tx = 2 * q.x, ty = 2 * q.y, tz = 2 * q.z;
  const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  const double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  R[0] = 1 - (tyy + tzz); R[3] = txy - twz;       R[6] = txz + twy;
  R[1] = txy + twz;       R[4] = 1 - (txx + tzz); R[7] = tyz - twx;
  R[2] = txz - twy;       R[5] = tyz + twx;       R[8] = 1 - (txx + tyy);
}

static double dot3(double a0, double a1, double a2, double b0, double b1, double b2) {
  return (a0 * b0 + a1 * b1) + a2 * b2;
}

/* C = A * B, all 3x3 column-major */
static void mat3_mul(const double* A, const double* B, double* C) {
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) C[c * 3 + r] = dot3(A[r], A[3 + r], A[6 + r], B[c * 3], B[c * 3 + 1], B[c * 3 + 2]);
}

/* ---------------------------------------------------------------------------------------------------
 * ActuatorLink / ActuatorChain (test/transform_chains.cc)
 * --------------------------------------------------------------------------------------------------- */
typedef struct {
  int links, n;
  quat q[MAX_LINKS];            /* parent_T_child.rotation */
  double t[MAX_LINKS][3];       /* parent_T_child.translation */
  unsigned char active[MAX_LINKS][6];
  double rotation_xyz[MAX_LINKS][3];
  /* cache, ActuatorChain::ShouldUpdate (transform_chains.cc:248-258) */
  int have_cache;
  double params_cached[3 * MAX_LINKS];
  /* outputs of Update */
  quat eff_q;
  double eff_t[3];
  double rot_D[3 * 6 * MAX_LINKS];   /* 3 x n column-major */
  double trans_D[3 * 6 * MAX_LINKS]; /* 3 x n column-major */
} chain_t;

static int active_count(const chain_t* c, int i) {
  int s = 0;
  for (int j = 0; j < 6; ++j) s += c->active[i][j] > 0;
  return s;
}
static int active_rot_count(const chain_t* c, int i) {
  return (c->active[i][0] > 0) + (c->active[i][1] > 0) + (c->active[i][2] > 0);
}

/* EulerAnglesFromSO3, so3.hpp:365-371 */
static void euler_from_quat(quat q, double* out) {
  out[0] = atan2(2 * (q.y * q.z + q.x * q.w), 1 - 2 * (q.x * q.x + q.y * q.y));
  out[1] = asin(fmax(fmin(1., -2 * (q.x * q.z - q.y * q.w)), -1.));
  out[2] = atan2(2 * (q.x * q.y + q.z * q.w), 1 - 2 * (q.y * q.y + q.z * q.z));
}

static int chain_init(chain_t* c, const mo_chain_ik_batch* b) {
  if (b->links > MAX_LINKS || b->links < 0) return -1;
  memset(c, 0, sizeof(*c));
  c->links = b->links;
  int total = 0;
  for (int i = 0; i < b->links; ++i) {
    const double* p = b->link_pose + 7 * i;
    c->q[i].w = p[0]; c->q[i].x = p[1]; c->q[i].y = p[2]; c->q[i].z = p[3];
    c->t[i][0] = p[4]; c->t[i][1] = p[5]; c->t[i][2] = p[6];
    for (int j = 0; j < 6; ++j) c->active[i][j] = b->link_mask[6 * i + j];
    total += active_count(c, i);
    if (active_rot_count(c, i) > 0) { /* ActuatorLink ctor, transform_chains.cc:107-122 */
      double e[3];
      euler_from_quat(quat_conj(c->q[i]), e);
      c->rotation_xyz[i][0] = -e[0]; c->rotation_xyz[i][1] = -e[1]; c->rotation_xyz[i][2] = -e[2];
    }
  }
  c->n = total;
  return total == b->n ? 0 : -1;
}

/* SO3FromEulerAngles_(xyz, XYZ): quaternion + right-tangent derivative wrt [x, y, z] (so3.hpp:293-338).
 * D is 3x3 column-major. */
static void so3_from_euler_xyz(const double* xyz, quat* q, double* D) {
  const double c_hx = cos(xyz[0] / 2), s_hx = sin(xyz[0] / 2);
  const double c_hy = cos(xyz[1] / 2), s_hy = sin(xyz[1] / 2);
  const double c_hz = cos(xyz[2] / 2), s_hz = sin(xyz[2] / 2);
  const double cy = 2 * c_hy * c_hy - 1, sy = 2 * c_hy * s_hy;
  const double cz = 2 * c_hz * c_hz - 1, sz = 2 * c_hz * s_hz;
  q->w = c_hx * c_hy * c_hz - s_hx * s_hy * s_hz;
  q->x = c_hx * s_hy * s_hz + c_hy * c_hz * s_hx;
  q->y = c_hx * c_hz * s_hy - c_hy * s_hx * s_hz;
  q->z = c_hx * c_hy * s_hz + c_hz * s_hx * s_hy;
  for (int i = 0; i < 9; ++i) D[i] = 0.0;
  D[0] = cy * cz; D[1] = -sz * cy; D[2] = sy; /* column 0 */
  D[3] = sz;      D[4] = cz;                  /* column 1 */
  D[8] = 1;                                   /* column 2 */
}

/* ActuatorLink::Compute, transform_chains.cc:125-158. J_out: 3 x ActiveRotationCount column-major. */
static void link_compute(const chain_t* c, int i, const double* params, int position, quat* q_out, double* t_out,
                         double* J_out) {
  if (active_rot_count(c, i) == 0) {
    t_out[0] = c->t[i][0]; t_out[1] = c->t[i][1]; t_out[2] = c->t[i][2];
    int param_index = position;
    for (int a = 0; a < 3; ++a)
      if (c->active[i][a + 3]) t_out[a] = params[param_index++];
    *q_out = c->q[i];
    return;
  }
  double pu[6] = {c->rotation_xyz[i][0], c->rotation_xyz[i][1], c->rotation_xyz[i][2], c->t[i][0], c->t[i][1], c->t[i][2]};
  int param_index = position;
  for (int a = 0; a < 6; ++a)
    if (c->active[i][a]) pu[a] = params[param_index++];
  double D[9];
  so3_from_euler_xyz(pu, q_out, D);
  int oi = 0;
  for (int axis = 0; axis < 3; ++axis)
    if (c->active[i][axis]) {
      J_out[3 * oi] = D[3 * axis]; J_out[3 * oi + 1] = D[3 * axis + 1]; J_out[3 * oi + 2] = D[3 * axis + 2];
      ++oi;
    }
  t_out[0] = pu[3]; t_out[1] = pu[4]; t_out[2] = pu[5];
}

/* ActuatorChain::Update (transform_chains.cc:165-244) with ComputeChain (:23-82) inlined. */
static void chain_update(chain_t* c, const double* params) {
  const int N = c->links, n = c->n;
  if (c->have_cache) { /* ShouldUpdate, :248-258 */
    int changed = 0;
    for (int i = 0; i < n; ++i)
      if (fabs(c->params_cached[i] - params[i]) > 1.0e-9) { changed = 1; break; }
    if (!changed) return;
  } else {
    for (int i = 0; i < 3 * n; ++i) { c->rot_D[i] = 0.0; c->trans_D[i] = 0.0; }
  }
  c->have_cache = 1;
  for (int i = 0; i < n; ++i) c->params_cached[i] = params[i];

  static _Thread_local quat lq[MAX_LINKS], i_R_end[MAX_LINKS + 1];
  static _Thread_local double lt[MAX_LINKS][3], i_t_end[MAX_LINKS + 1][3];
  static _Thread_local double tDt[MAX_LINKS][9], tDr[MAX_LINKS][9], rDr[MAX_LINKS][9];
  for (int i = 0, position = 0; i < N; ++i) {
    link_compute(c, i, params, position, &lq[i], lt[i], c->rot_D + 3 * position);
    position += active_count(c, i);
  }
  if (N == 0) return;
  /* backwards rotations / translations (:36-52) */
  i_R_end[N].w = 1; i_R_end[N].x = 0; i_R_end[N].y = 0; i_R_end[N].z = 0;
  for (int i = N - 1; i >= 0; --i) i_R_end[i] = quat_mul(lq[i], i_R_end[i + 1]);
  i_t_end[N][0] = i_t_end[N][1] = i_t_end[N][2] = 0.0;
  for (int i = N - 1; i >= 0; --i) {
    double r[3];
    quat_rotate(lq[i], i_t_end[i + 1], r);
    i_t_end[i][0] = r[0] + lt[i][0]; i_t_end[i][1] = r[1] + lt[i][1]; i_t_end[i][2] = r[2] + lt[i][2];
  }
  /* translation_D_translation (:56-61) */
  quat start_R_i = {1, 0, 0, 0};
  for (int i = 0; i < N; ++i) {
    quat_matrix(start_R_i, tDt[i]);
    start_R_i = quat_mul(start_R_i, lq[i]);
  }
  /* translation_D_rotation (:66-73): start_R_[i+1] * Skew3(-[i+1]_t_N) */
  for (int i = 0; i < N - 1; ++i) {
    const double v[3] = {-i_t_end[i + 1][0], -i_t_end[i + 1][1], -i_t_end[i + 1][2]};
    const double S[9] = {0, v[2], -v[1], -v[2], 0, v[0], v[1], -v[0], 0}; /* column-major Skew3(v) */
    mat3_mul(tDt[i + 1], S, tDr[i]);
  }
  for (int k = 0; k < 9; ++k) tDr[N - 1][k] = 0.0;
  /* rotation_D_rotation (:76-81) */
  for (int i = 0; i < N - 1; ++i) quat_matrix(quat_conj(i_R_end[i + 1]), rDr[i]);
  for (int k = 0; k < 9; ++k) rDr[N - 1][k] = (k % 4 == 0) ? 1.0 : 0.0;

  /* chain rule onto the active parameters (:198-243) */
  for (int i = 0, position = 0; i < N; ++i) {
    const int ac = active_count(c, i), nr = active_rot_count(c, i);
    if (ac == 0) continue;
    double* rot_D_angles = c->rot_D + 3 * position;
    double* trans_D = c->trans_D + 3 * position;
    for (int a = 0; a < nr; ++a) {
      const double* col = rot_D_angles + 3 * a;
      for (int r = 0; r < 3; ++r) trans_D[3 * a + r] = dot3(tDr[i][r], tDr[i][3 + r], tDr[i][6 + r], col[0], col[1], col[2]);
    }
    for (int a = 0; a < nr; ++a) {
      double* col = rot_D_angles + 3 * a;
      double tmp[3];
      for (int r = 0; r < 3; ++r) tmp[r] = dot3(rDr[i][r], rDr[i][3 + r], rDr[i][6 + r], col[0], col[1], col[2]);
      col[0] = tmp[0]; col[1] = tmp[1]; col[2] = tmp[2];
    }
    if (ac - nr > 0) {
      int oi = 0;
      for (int axis = 0; axis < 3; ++axis)
        if (c->active[i][axis + 3]) {
          for (int r = 0; r < 3; ++r) trans_D[3 * (nr + oi) + r] = tDt[i][3 * axis + r];
          ++oi;
        }
    }
    position += ac;
  }
  c->eff_q = i_R_end[0];
  c->eff_t[0] = i_t_end[0][0]; c->eff_t[1] = i_t_end[0][1]; c->eff_t[2] = i_t_end[0][2];
}

/* ActuatorChain::Update for `batch` parameter vectors, a fresh chain (empty cache) per problem. */
int oracle_chain_fk_batch(const mo_chain_ik_batch* b, const double* params, double* translation, double* rotation_q,
                          double* t_D_params, double* r_D_params) {
  if (!b || !params) return MO_ERR_INVALID_ARGUMENT;
  chain_t* c = (chain_t*)malloc(sizeof(chain_t));
  const int n = b->n;
  for (int64_t p = 0; p < b->batch; ++p) {
    if (chain_init(c, b)) { free(c); return MO_ERR_INVALID_ARGUMENT; }
    chain_update(c, params + p * n);
    if (translation) memcpy(translation + 3 * p, c->eff_t, 3 * sizeof(double));
    if (rotation_q) {
      rotation_q[4 * p] = c->eff_q.w; rotation_q[4 * p + 1] = c->eff_q.x;
      rotation_q[4 * p + 2] = c->eff_q.y; rotation_q[4 * p + 3] = c->eff_q.z;
    }
    if (t_D_params) memcpy(t_D_params + 3 * n * p, c->trans_D, 3 * n * sizeof(double));
    if (r_D_params) memcpy(r_D_params + 3 * n * p, c->rot_D, 3 * n * sizeof(double));
  }
  free(c);
  return MO_OK;
}

/* ---------------------------------------------------------------------------------------------------
 * Residual::Model::UpdateHessian (residual.hpp:186-226) for a dense R x n block with identity index:
 * lower triangle += J_i . J_j, c += J_i . r; returns 1/2 |r|^2. J column-major with leading dimension ld.
 * --------------------------------------------------------------------------------------------------- */
static double update_hessian(int R, int n, const double* J, int ld, const double* r, double* G, double* c) {
  for (int row = 0; row < n; ++row) {
    for (int col = 0; col <= row; ++col) {
      double d = 0.0;
      for (int i = 0; i < R; ++i) d += J[row * ld + i] * J[col * ld + i];
      G[col * n + row] += d;
    }
    double d = 0.0;
    for (int i = 0; i < R; ++i) d += J[row * ld + i] * r[i];
    c[row] += d;
  }
  double s = 0.0;
  for (int i = 0; i < R; ++i) s += r[i] * r[i];
  return 0.5 * s;
}

/* LinearizeAndFillQP, nonlinear.cc:170-214, dense Jacobians. G, A column-major. */
static void linearize_dense(int n, int r_cost, int cost_block, int k, int m, const double* x, const double* Jc,
                            const double* rc, const double* Je, const double* re, double lambda,
                            const int32_t* ivar, const double* ia, const double* ib, double* G, double* c, double* A,
                            double* b, double* ib_shift, double* errors) {
  const int cb = cost_block > 0 ? cost_block : r_cost;
  for (int i = 0; i < n * n; ++i) G[i] = 0.0;
  for (int i = 0; i < n; ++i) c[i] = 0.0;
  double f = 0.0;
  for (int row = 0; row < r_cost; row += cb) f += update_hessian(cb, n, Jc + row, r_cost, rc + row, G, c);
  if (lambda > 0)
    for (int i = 0; i < n; ++i) G[i * n + i] += lambda;
  double eq = 0.0;
  for (int row = 0; row < k; ++row) { /* one 1-row Residual per equality row: UpdateJacobian, residual.hpp:230-250 */
    b[row] = re[row];
    for (int col = 0; col < n; ++col) A[col * k + row] = Je[col * k + row];
    eq += fabs(b[row]);
  }
  for (int i = 0; i < m; ++i) ib_shift[i] = ia[i] * x[ivar[i]] + ib[i]; /* ShiftTo, qp.hpp:57-59 */
  errors[0] = f;
  errors[1] = eq;
}

int oracle_nls_linearize_batch(const mo_nls_linearize_args* a) {
  if (!a || a->n <= 0 || a->r_cost < 0 || a->r_eq < 0 || a->m < 0) return MO_ERR_INVALID_ARGUMENT;
  const int n = a->n, k = a->r_eq, m = a->m;
  for (int64_t p = 0; p < a->batch; ++p)
    linearize_dense(n, a->r_cost, a->cost_block, k, m, a->x + p * n, a->J_cost + p * a->r_cost * n,
                    a->r_cost_v + p * a->r_cost, k ? a->J_eq + p * k * n : NULL, k ? a->r_eq_v + p * k : NULL,
                    a->lambda[p], a->ineq_var + p * m, a->ineq_a + p * m, a->ineq_b + p * m, a->G + p * n * n,
                    a->c + p * n, k ? a->A_eq + p * k * n : NULL, k ? a->b_eq + p * k : NULL,
                    a->ineq_b_shifted + p * m, a->errors + 2 * p);
  return MO_OK;
}

/* ---------------------------------------------------------------------------------------------------
 * Chain residuals: row = w . root_t_effector - target, J row = w' * translation_D_params
 * --------------------------------------------------------------------------------------------------- */
static void chain_rows(chain_t* ch, const double* x, int rows, const double* w, const double* target, double* r,
                       double* J /* rows x n column-major, or NULL */) {
  chain_update(ch, x);
  for (int i = 0; i < rows; ++i) {
    const double* wi = w + 3 * i;
    r[i] = dot3(wi[0], wi[1], wi[2], ch->eff_t[0], ch->eff_t[1], ch->eff_t[2]) - target[i];
    if (J)
      for (int c = 0; c < ch->n; ++c)
        J[c * rows + i] = dot3(wi[0], wi[1], wi[2], ch->trans_D[3 * c], ch->trans_D[3 * c + 1], ch->trans_D[3 * c + 2]);
  }
}

/* EvaluateNonlinearErrors, nonlinear.cc:279-293 */
static void eval_errors(chain_t* ch, const mo_chain_ik_batch* b, int64_t p, const double* x, double* errors) {
  double r[16];
  const int cb = b->cost_block > 0 ? b->cost_block : b->n_cost;
  double f = 0.0, eq = 0.0;
  chain_rows(ch, x, b->n_cost, b->cost_w, b->cost_target + p * b->n_cost, r, NULL);
  for (int row = 0; row < b->n_cost; row += cb) {
    double s = 0.0;
    for (int i = 0; i < cb; ++i) s += r[row + i] * r[row + i];
    f += 0.5 * s;
  }
  if (b->n_eq) {
    chain_rows(ch, x, b->n_eq, b->eq_w, b->eq_target + p * b->n_eq, r, NULL);
    for (int i = 0; i < b->n_eq; ++i) eq += fabs(r[i]);
  }
  errors[0] = f;
  errors[1] = eq;
}

static double mod_pi(double angle) { /* angle_utils.hpp:9-17 */
  const double pi = M_PI, two_pi = 2 * M_PI;
  angle = fmod(angle, two_pi);
  angle += (angle < 0) * two_pi;
  angle -= (angle > pi) * two_pi;
  return angle;
}

static double sign(double x) { return x > 0 ? 1.0 : (x < 0 ? -1.0 : 0.0); }

/* QuadraticApproxMinimum / CubicApproxCoeffs / CubicApproxMinimum, nonlinear.cc:531-605. Return 0 = nullopt. */
static int quadratic_approx_minimum(double phi_0, double phi_prime_0, double alpha_0, double phi_alpha_0, double* out) {
  const double numerator = phi_alpha_0 - phi_prime_0 * alpha_0 - phi_0;
  if (phi_prime_0 > 0 || numerator <= 0) return 0;
  *out = -phi_prime_0 * alpha_0 * alpha_0 / (2.0 * numerator);
  return 1;
}

static void cubic_approx_coeffs(double phi_0, double phi_prime_0, double alpha_0, double phi_alpha_0, double alpha_1,
                                double phi_alpha_1, double* ab) {
  /* A = [[a0^3, a0^2],[a1^3, a1^2]]; ab = A.inverse() * rhs (Eigen 2x2 inverse: adjugate / determinant) */
  const double a = alpha_0 * alpha_0 * alpha_0, bb = alpha_0 * alpha_0;
  const double c = alpha_1 * alpha_1 * alpha_1, d = alpha_1 * alpha_1;
  const double rhs0 = phi_alpha_0 - phi_0 - phi_prime_0 * alpha_0;
  const double rhs1 = phi_alpha_1 - phi_0 - phi_prime_0 * alpha_1;
  const double invdet = 1.0 / (a * d - bb * c);
  const double i00 = d * invdet, i01 = -bb * invdet, i10 = -c * invdet, i11 = a * invdet;
  ab[0] = i00 * rhs0 + i01 * rhs1;
  ab[1] = i10 * rhs0 + i11 * rhs1;
}

static int cubic_approx_minimum(double phi_prime_0, const double* ab, double* out) {
  const double a = ab[0], b = ab[1];
  const double arg_sqrt = b * b - 3 * a * phi_prime_0;
  if (a == 0.0 || arg_sqrt < -1.0e-12) return 0;
  *out = (-b + sqrt(fmax(arg_sqrt, 0.))) / (3 * a);
  return 1;
}

/* ---------------------------------------------------------------------------------------------------
 * ConstrainedNonlinearLeastSquares::Solve (nonlinear.cc:75-158) for one chain problem
 * --------------------------------------------------------------------------------------------------- */
typedef struct {
  int n, k, m;
  double *G, *c, *A, *b, *ib_shift, *Jc, *Je, *x, *cand, *dx, *vars;
  chain_t* chain;
} nls_work;

static double total_err(const double* e, double penalty) { return e[0] + penalty * e[1]; }

static int check_nls_params(const mo_nls_params* p) { /* CheckParams, nonlinear.cc:49-73 */
  return p->max_iterations >= 0 && p->max_qp_iterations >= 1 && p->termination_kkt_tolerance > 0 &&
         p->absolute_exit_tol > 0 && p->max_line_search_iterations >= 0 && p->relative_exit_tol >= 0 &&
         p->relative_exit_tol <= 1 && p->absolute_first_derivative_tol >= 0 && p->armijo_search_tau > 0 &&
         p->armijo_search_tau < 1 && p->equality_penalty_initial >= 0 && p->equality_penalty_scale_factor >= 1.0 &&
         p->equality_penalty_rho >= 0 && p->equality_penalty_rho < 1 && p->max_lambda >= 0 &&
         p->min_lambda <= p->max_lambda && p->lambda_initial >= p->min_lambda && p->lambda_initial <= p->max_lambda &&
         p->lambda_failure_init >= 0 && p->lambda_decrease_on_success >= 0 && p->lambda_decrease_on_success < 1.0 &&
         p->lambda_decrease_on_restore >= 0 && p->lambda_decrease_on_restore < 1.0;
}

static void retract(const nls_work* w, const mo_nls_params* prm, double alpha) { /* RetractCandidateVars, :160-168 */
  for (int i = 0; i < w->n; ++i) {
    if (prm->retraction == MO_RETRACT_MOD_PI) w->cand[i] = mod_pi(w->x[i] + w->dx[i] * alpha);
    else w->cand[i] = w->x[i] + w->dx[i] * alpha;
  }
}

static void solve_one(const mo_chain_ik_batch* b, const mo_nls_params* prm, int64_t p, nls_work* w, int* term_out,
                      int* iters_out, int* qp_iters_out, int* ls_out) {
  const int n = w->n, k = w->k, m = w->m;
  chain_t* ch = w->chain;
  ch->have_cache = 0;
  for (int i = 0; i < n; ++i) w->x[i] = b->x0[p * n + i];
  int state = 0; /* OptimizerState: 0 NOMINAL, 1 ATTEMPTING_RESTORE_LM */
  double lambda = prm->lambda_initial, penalty = prm->equality_penalty_initial;
  int term = MO_NLS_MAX_ITERATIONS, iters = 0, qp_total = 0, ls_total = 0;
  double rc[16], re[16];
  for (int iter = 0; iter < prm->max_iterations; ++iter) {
    /* LinearizeAndFillQP */
    double errors_pre[2];
    chain_rows(ch, w->x, b->n_cost, b->cost_w, b->cost_target + p * b->n_cost, rc, w->Jc);
    if (k) chain_rows(ch, w->x, k, b->eq_w, b->eq_target + p * k, re, w->Je);
    linearize_dense(n, b->n_cost, b->cost_block, k, m, w->x, w->Jc, rc, w->Je, re, lambda, b->ineq_var, b->ineq_a,
                    b->ineq_b, w->G, w->c, w->A, w->b, w->ib_shift, errors_pre);
    /* ComputeStepDirection, :216-259 */
    mo_qp_batch qb;
    memset(&qb, 0, sizeof(qb));
    qb.n = n; qb.k = k; qb.m = m; qb.batch = 1;
    qb.G = w->G; qb.c = w->c; qb.A_eq = w->A; qb.b_eq = w->b;
    qb.ineq_var = b->ineq_var; qb.ineq_a = b->ineq_a; qb.ineq_b = w->ib_shift;
    mo_qp_params qp;
    qp.initial_mu = 1.0; qp.sigma = 0.1; qp.termination_kkt_tol = prm->termination_kkt_tolerance;
    qp.termination_complementarity_tol = 1.0e-6; qp.max_iterations = prm->max_qp_iterations;
    qp.barrier_strategy = MO_BARRIER_COMPLEMENTARITY; qp.decrease_mu_only_on_small_error = 0;
    qp.initial_guess_method = k ? MO_GUESS_SOLVE_EQUALITY_CONSTRAINED : MO_GUESS_NAIVE;
    qp.initialize_mu_with_complementarity = 0; qp.reserved = 0;
    int32_t qstatus = -1, qiters = 0;
    double lag[2] = {0, 0};
    const int null_space = m == 0 && k > 0; /* solver choice, nonlinear.cc:83-89 */
    if (null_space) {
      const int ns = oracle_qp_nullspace_one(n, k, w->G, w->c, w->A, w->b, w->dx); /* dx = 0 unless SUCCESS, :248-258 */
      if (ns == 1) { term = MO_NLS_QP_INDEFINITE; break; } /* QPWasIndefinite, nonlinear.cc:103-105 */
      if (ns != 0) { term = MO_NLS_QP_FAILED; break; }
      qp_total += 1; /* NumQPIterations counts one per null-space solve, structs.cc:208-219 */
    } else {
      mo_qp_out qo;
      memset(&qo, 0, sizeof(qo));
      qo.vars = w->vars; qo.status = &qstatus; qo.iters = &qiters; qo.lagrange = lag;
      oracle_qp_solve_batch(&qb, &qp, &qo);
      if (qstatus >= MO_QP_FAILED_FACTORIZATION) { term = MO_NLS_QP_FAILED; break; } /* the reference throws */
      qp_total += qiters;
      for (int i = 0; i < n; ++i) w->dx[i] = w->vars[i];
    }
    /* ComputeQPCostDerivative, :457-482 */
    double d_f = 0.0, d_eq = 0.0;
    for (int i = 0; i < n; ++i) d_f += w->c[i] * w->dx[i];
    for (int r = 0; r < k; ++r) {
      double d = 0.0;
      for (int i = 0; i < n; ++i) d += w->A[i * k + r] * w->dx[i];
      d_eq += sign(w->b[r]) * d;
    }
    /* SelectPenalty, :484-502 (interior point branch: |y|_inf) */
    if (k) {
      double new_penalty = lag[1];
      if (null_space) { /* no multipliers: inequality (18.36), nonlinear.cc:492-500 */
        double l1 = 0.0, cdx = 0.0, quad = 0.0;
        for (int r = 0; r < k; ++r) l1 += fabs(w->b[r]);
        if (l1 < DBL_EPSILON) l1 = DBL_EPSILON;
        for (int i = 0; i < n; ++i) cdx += w->c[i] * w->dx[i];
        for (int i = 0; i < n; ++i) {
          double acc = 0.0;
          for (int j = 0; j < n; ++j) acc += (i >= j ? w->G[(size_t)j * n + i] : w->G[(size_t)i * n + j]) * w->dx[j];
          quad += w->dx[i] * acc;
        }
        new_penalty = (cdx + 0.5 * (quad > 0.0 ? quad : 0.0)) / ((1 - prm->equality_penalty_rho) * l1);
      }
      if (new_penalty > penalty) penalty = new_penalty * prm->equality_penalty_scale_factor;
    }
    /* SelectStepSize, :351-416 */
    const double armijo_c1 = 1.0e-4;
    const double dd_total = d_f + penalty * d_eq;
    const double dd_linf = fmax(fabs(d_f), fabs(d_eq));
    int step_result = 1; /* 0 SUCCESS 1 MAX_ITERATIONS 2 FIRST_ORDER 3 POSITIVE_DERIVATIVE 4 NON_FINITE 5 INVALID_ALPHA */
    double alpha = 1.0, step_alpha[64], step_err[64][2];
    int nsteps = 0;
    for (int it = 0; it <= prm->max_line_search_iterations; ++it) {
      if (prm->line_search_strategy == MO_LS_POLYNOMIAL_APPROXIMATION) {
        if (it > 0) {
          double na = 0.0;
          int has;
          const double phi0 = total_err(errors_pre, penalty);
          if (it == 1) {
            has = quadratic_approx_minimum(phi0, dd_total, step_alpha[nsteps - 1], total_err(step_err[nsteps - 1], penalty), &na);
          } else {
            double ab[2];
            cubic_approx_coeffs(phi0, dd_total, step_alpha[nsteps - 2], total_err(step_err[nsteps - 2], penalty),
                                step_alpha[nsteps - 1], total_err(step_err[nsteps - 1], penalty), ab);
            has = cubic_approx_minimum(dd_total, ab, &na);
          }
          if (!has || !isfinite(na) || na <= 0.0 || na >= alpha) { step_result = 5; break; }
          alpha = na;
        }
      } else if (it > 0) {
        alpha = alpha * prm->armijo_search_tau;
      }
      retract(w, prm, alpha);
      double es[2];
      eval_errors(ch, b, p, w->cand, es);
      if (nsteps < 64) { step_alpha[nsteps] = alpha; step_err[nsteps][0] = es[0]; step_err[nsteps][1] = es[1]; ++nsteps; }
      ++ls_total;
      if (!isfinite(es[0]) || !isfinite(es[1])) { step_result = 4; break; }
      if (dd_linf < prm->absolute_first_derivative_tol) { step_result = 2; break; }
      else if (dd_total > 0) { step_result = 3; break; }
      if (total_err(es, penalty) <= total_err(errors_pre, penalty) + dd_total * alpha * armijo_c1) { step_result = 0; break; }
    }
    /* UpdateLambdaAndCheckExitConditions, :296-343 */
    int exit_state = -1;
    if (step_result == 0) {
      for (int i = 0; i < n; ++i) { const double t = w->x[i]; w->x[i] = w->cand[i]; w->cand[i] = t; }
      if (state == 1) lambda = fmax(lambda * prm->lambda_decrease_on_restore, prm->min_lambda);
      else lambda = fmax(lambda * prm->lambda_decrease_on_success, prm->min_lambda);
      state = 0;
      const double* fe = step_err[nsteps - 1];
      if (fmax(fe[0], fe[1]) < prm->absolute_exit_tol) exit_state = MO_NLS_SATISFIED_ABSOLUTE_TOL;
      else if (total_err(fe, penalty) > total_err(errors_pre, penalty) * (1 - prm->relative_exit_tol))
        exit_state = MO_NLS_SATISFIED_RELATIVE_TOL;
    } else if (step_result == 2) {
      exit_state = MO_NLS_SATISFIED_FIRST_ORDER_TOL;
    } else if (step_result == 1 || step_result == 3) {
      if (state == 0) { lambda = fmax(prm->lambda_failure_init, lambda * 10.0); state = 1; }
      else lambda *= 10.0;
      if (lambda > prm->max_lambda) exit_state = MO_NLS_MAX_LAMBDA;
    }
    ++iters;
    if (exit_state >= 0) { term = exit_state; break; }
  }
  *term_out = term; *iters_out = iters; *qp_iters_out = qp_total; *ls_out = ls_total;
}

int64_t oracle_nls_solve_chain_ik_batch(const mo_chain_ik_batch* b, const mo_nls_params* prm, const mo_nls_out* out) {
  if (!b || !prm || !out || !out->x || !out->termination || !out->iterations) return MO_ERR_INVALID_ARGUMENT;
  if (!check_nls_params(prm)) return MO_ERR_INVALID_PARAMS;
  if (b->n_cost > 16 || b->n_eq > 16 || b->n <= 0) return MO_ERR_INVALID_ARGUMENT;
  const int n = b->n, k = b->n_eq, m = b->m;
  nls_work w;
  w.n = n; w.k = k; w.m = m;
  w.chain = (chain_t*)malloc(sizeof(chain_t));
  if (chain_init(w.chain, b)) { free(w.chain); return MO_ERR_INVALID_ARGUMENT; }
  w.G = (double*)calloc((size_t)n * n + 1, 8); w.c = (double*)calloc(n + 1, 8);
  w.A = (double*)calloc((size_t)k * n + 1, 8); w.b = (double*)calloc(k + 1, 8);
  w.ib_shift = (double*)calloc(m + 1, 8); w.Jc = (double*)calloc((size_t)b->n_cost * n + 1, 8);
  w.Je = (double*)calloc((size_t)k * n + 1, 8); w.x = (double*)calloc(n + 1, 8); w.cand = (double*)calloc(n + 1, 8);
  w.dx = (double*)calloc(n + 1, 8); w.vars = (double*)calloc((size_t)n + 2 * m + k + 1, 8);
  int64_t total = 0;
  for (int64_t p = 0; p < b->batch; ++p) {
    int term, iters, qpi, ls;
    solve_one(b, prm, p, &w, &term, &iters, &qpi, &ls);
    for (int i = 0; i < n; ++i) out->x[p * n + i] = w.x[i];
    out->termination[p] = term; out->iterations[p] = iters;
    if (out->qp_iterations) out->qp_iterations[p] = qpi;
    if (out->ls_steps) out->ls_steps[p] = ls;
    if (out->errors) eval_errors(w.chain, b, p, w.x, out->errors + 2 * p);
    total += qpi;
  }
  free(w.chain); free(w.G); free(w.c); free(w.A); free(w.b); free(w.ib_shift); free(w.Jc); free(w.Je); free(w.x);
  free(w.cand); free(w.dx); free(w.vars);
  return total;
}
