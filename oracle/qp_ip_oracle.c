/*
 * TEST INFRASTRUCTURE — not part of the product. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product path never does.
 *
 * Plain-C restatement of the reference's interior point QP solver (gareth-cross/mini_opt,
 * source/qp.cc) INCLUDING the parts of Eigen 3.4.0 it leans on (LDLT<MatrixXd,Lower> with diagonal
 * pivoting and the explicit inverse through ldlt.solveInPlace(I)). It follows the reference's
 * algorithm step for step — deliberately not the CUDA path's algorithm (which factors without pivoting
 * and never forms an inverse) — so that agreement between the two is evidence, not tautology.
 *
 * Parity: PINNED. tests/test_oracle_vs_reference.py checks this file against the reference itself
 * (oracle/_ref/libminiopt_ref.so, built from /root/reference by oracle/Makefile) and against the golden
 * vectors under tests/golden/ produced by the reference's own generator and solver.
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "mini_opt_b200.h"

typedef struct {
  int n, k, m; /* m = number of constraints actually present for this problem */
  const double* G;
  const double* c;
  const double* A; /* k x n column-major */
  const double* b;
  const int32_t* var;
  const double* ia;
  const double* ib;
  int box_pairs;
} oq_problem;

typedef struct {
  int p;         /* n + k */
  double* H;     /* p x p column-major, zeroed once (Setup, qp.cc:46-47) */
  double* LD;    /* factor storage (Eigen LDLT::m_matrix) */
  double* Hinv;  /* explicit inverse (qp.cc:310-311) */
  int* transp;   /* LDLT::m_transpositions */
  double* tmp;   /* LDLT::m_temporary */
  double *vars, *r, *delta, *delta_aff, *r_aug; /* [x|s|y|z] layout, qp.cc:548-582 */
} oq_work;

static int con_var(const oq_problem* q, int i) { return q->box_pairs ? i / 2 : q->var[i]; }
static double con_a(const oq_problem* q, int i) { return q->box_pairs ? ((i % 2 == 0) ? 1.0 : -1.0) : q->ia[i]; }

#define XB(w, q) ((w))
#define SB(w, q) ((w) + (q)->n)
#define YB(w, q) ((w) + (q)->n + (q)->m)
#define ZB(w, q) ((w) + (q)->n + (q)->m + (q)->k)

/* Setup, qp.cc:20-73 */
static void work_alloc(oq_work* w, const oq_problem* q) {
  const int p = q->n + q->k, t = q->n + 2 * q->m + q->k;
  w->p = p;
  w->H = (double*)calloc((size_t)(p * p + 1), sizeof(double));
  w->LD = (double*)calloc((size_t)(p * p + 1), sizeof(double));
  w->Hinv = (double*)calloc((size_t)(p * p + 1), sizeof(double));
  w->transp = (int*)calloc((size_t)(p + 1), sizeof(int));
  w->tmp = (double*)calloc((size_t)(p + 1), sizeof(double));
  w->vars = (double*)calloc((size_t)(t + 1), sizeof(double));
  w->r = (double*)calloc((size_t)(t + 1), sizeof(double));
  w->delta = (double*)calloc((size_t)(t + 1), sizeof(double));
  w->delta_aff = (double*)calloc((size_t)(t + 1), sizeof(double));
  w->r_aug = (double*)calloc((size_t)(q->n + 1), sizeof(double));
}

static void work_free(oq_work* w) {
  free(w->H); free(w->LD); free(w->Hinv); free(w->transp); free(w->tmp);
  free(w->vars); free(w->r); free(w->delta); free(w->delta_aff); free(w->r_aug);
}

/*
 * Eigen 3.4.0 internal::ldlt_inplace<Lower>::unblocked, Eigen/src/Cholesky/LDLT.h:297-403.
 * In-place LDL^T with diagonal pivoting on the lower triangle of `mat` (size x size, column-major).
 * Returns 1 on success (Eigen's `ret`), 0 for NumericalIssue.
 */
static int eigen_ldlt_lower_unblocked(double* mat, int size, int* transp, double* temp) {
#define M(i, j) mat[(size_t)(j) * size + (i)]
  int found_zero_pivot = 0, ret = 1;
  if (size <= 1) { /* LDLT.h:311-319 */
    if (size == 1) transp[0] = 0;
    return 1;
  }
  for (int k = 0; k < size; ++k) {
    /* largest |diagonal| of the trailing block, first one wins ties (LDLT.h:324-326) */
    int big = k;
    double best = fabs(M(k, k));
    for (int i = k + 1; i < size; ++i) {
      if (fabs(M(i, i)) > best) { best = fabs(M(i, i)); big = i; }
    }
    transp[k] = big;
    if (k != big) { /* symmetric row/column swap restricted to the lower triangle (LDLT.h:329-345) */
      const int s = size - big - 1;
      for (int j = 0; j < k; ++j) { double t = M(k, j); M(k, j) = M(big, j); M(big, j) = t; }
      for (int i = 0; i < s; ++i) { double t = M(big + 1 + i, k); M(big + 1 + i, k) = M(big + 1 + i, big); M(big + 1 + i, big) = t; }
      { double t = M(k, k); M(k, k) = M(big, big); M(big, big) = t; }
      for (int i = k + 1; i < big; ++i) { double t = M(i, k); M(i, k) = M(big, i); M(big, i) = t; }
    }
    const int rs = size - k - 1;
    if (k > 0) { /* LDLT.h:356-362 */
      double dot = 0.0;
      for (int j = 0; j < k; ++j) { temp[j] = M(j, j) * M(k, j); dot += M(k, j) * temp[j]; }
      M(k, k) -= dot;
      for (int i = 0; i < rs; ++i) {
        double acc = 0.0;
        for (int j = 0; j < k; ++j) acc += M(k + 1 + i, j) * temp[j];
        M(k + 1 + i, k) -= acc;
      }
    }
    const double akk = M(k, k);
    const int pivot_is_valid = fabs(akk) > 0.0;
    if (k == 0 && !pivot_is_valid) { /* whole diagonal zero (LDLT.h:371-382) */
      for (int j = 0; j < size; ++j) {
        transp[j] = j;
        for (int i = j + 1; i < size; ++i) ret = ret && (M(i, j) == 0.0);
      }
      return ret;
    }
    if (rs > 0 && pivot_is_valid) {
      for (int i = 0; i < rs; ++i) M(k + 1 + i, k) /= akk; /* LDLT.h:384-385 */
    } else if (rs > 0) {
      for (int i = 0; i < rs; ++i) ret = ret && (M(k + 1 + i, k) == 0.0);
    }
    if (found_zero_pivot && pivot_is_valid) ret = 0; /* LDLT.h:389-390 */
    else if (!pivot_is_valid) found_zero_pivot = 1;
  }
  return ret;
#undef M
}

/* Eigen LDLT::_solve_impl_transposed, LDLT.h:574-611, for `ncols` right-hand sides stored column-major in X. */
static void eigen_ldlt_solve_inplace(const double* mat, const int* transp, int size, double* X, int ncols) {
#define M(i, j) mat[(size_t)(j) * size + (i)]
  for (int col = 0; col < ncols; ++col) {
    double* x = X + (size_t)col * size;
    for (int i = 0; i < size; ++i) { /* dst = P b */
      const int t = transp[i];
      if (t != i) { double v = x[i]; x[i] = x[t]; x[t] = v; }
    }
    for (int j = 0; j < size; ++j) /* unit-lower forward substitution */
      for (int i = j + 1; i < size; ++i) x[i] -= M(i, j) * x[j];
    for (int i = 0; i < size; ++i) { /* pseudo-inverse of D with tolerance DBL_MIN (LDLT.h:595-602) */
      if (fabs(M(i, i)) > DBL_MIN) x[i] /= M(i, i);
      else x[i] = 0.0;
    }
    for (int i = size - 1; i >= 0; --i) { /* L^T back substitution */
      double acc = x[i];
      for (int j = i + 1; j < size; ++j) acc -= M(j, i) * x[j];
      x[i] = acc;
    }
    for (int i = size - 1; i >= 0; --i) { /* dst = P^T dst */
      const int t = transp[i];
      if (t != i) { double v = x[i]; x[i] = x[t]; x[t] = v; }
    }
  }
#undef M
}

/* EvaluateKKTConditions, qp.cc:391-420 */
static void evaluate_kkt(const oq_problem* q, oq_work* w, int include_inequalities) {
  const int n = q->n, k = q->k, m = q->m;
  const double *x = XB(w->vars, q), *s = SB(w->vars, q), *y = YB(w->vars, q), *z = ZB(w->vars, q);
  double *r_d = XB(w->r, q), *r_comp = SB(w->r, q), *r_pe = YB(w->r, q), *r_pi = ZB(w->r, q);
  for (int i = 0; i < n; ++i) { /* G.selfadjointView<Lower>() * x + c */
    double acc = 0.0;
    for (int j = 0; j < n; ++j) acc += (i >= j ? q->G[(size_t)j * n + i] : q->G[(size_t)i * n + j]) * x[j];
    r_d[i] = acc + q->c[i];
  }
  if (k > 0) {
    for (int i = 0; i < n; ++i) {
      double acc = 0.0;
      for (int r = 0; r < k; ++r) acc += q->A[(size_t)i * k + r] * y[r];
      r_d[i] -= acc;
    }
    for (int r = 0; r < k; ++r) {
      double acc = 0.0;
      for (int j = 0; j < n; ++j) acc += q->A[(size_t)j * k + r] * x[j];
      r_pe[r] = acc + q->b[r];
    }
  }
  if (include_inequalities) {
    for (int i = 0; i < m; ++i) {
      const int v = con_var(q, i);
      const double a = con_a(q, i);
      r_d[v] -= a * z[i];
      r_pi[i] = a * x[v] + q->ib[i] - s[i];
      r_comp[i] = s[i] * z[i];
    }
  }
}

/* ComputeLDLT, qp.cc:275-316. Returns 0 ok, MO_QP_NONPOSITIVE_SLACK or MO_QP_FAILED_FACTORIZATION. */
static int compute_ldlt(const oq_problem* q, oq_work* w, int include_inequalities) {
  const int n = q->n, k = q->k, m = q->m, p = w->p;
  const double *s = SB(w->vars, q), *z = ZB(w->vars, q);
  if (include_inequalities)
    for (int i = 0; i < m; ++i)
      if (!(s[i] > 0.0)) return MO_QP_NONPOSITIVE_SLACK; /* F_ASSERT, qp.cc:285 */
  for (int j = 0; j < n; ++j)
    for (int i = j; i < n; ++i) w->H[(size_t)j * p + i] = q->G[(size_t)j * n + i];
  for (int j = 0; j < n; ++j)
    for (int r = 0; r < k; ++r) w->H[(size_t)j * p + n + r] = q->A[(size_t)j * k + r];
  if (include_inequalities) {
    for (int i = 0; i < m; ++i) {
      const int v = con_var(q, i);
      const double a = con_a(q, i);
      w->H[(size_t)v * p + v] += a * (z[i] / s[i]) * a;
    }
  }
  memcpy(w->LD, w->H, sizeof(double) * (size_t)p * p);
  if (!eigen_ldlt_lower_unblocked(w->LD, p, w->transp, w->tmp)) return MO_QP_FAILED_FACTORIZATION;
  memset(w->Hinv, 0, sizeof(double) * (size_t)p * p);
  for (int i = 0; i < p; ++i) w->Hinv[(size_t)i * p + i] = 1.0;
  eigen_ldlt_solve_inplace(w->LD, w->transp, p, w->Hinv, p);
  const int t = q->n + 2 * q->m + q->k;
  memset(w->delta, 0, sizeof(double) * t);
  memset(w->delta_aff, 0, sizeof(double) * t);
  return 0;
}

/* DIAGNOSTIC modes (tools/diagnose_status.py; default 0 = the reference's algorithm). They restate, on the CPU, the two
 * ways the CUDA kernels solve the Newton system, so that a status difference against the reference can be attributed:
 *   1  Eigen-pivoted LDLT (as the reference) but ONE substitution solve of the right-hand side, no explicit inverse
 *      (the generic kernel, qp_generic.cuh)
 *   2  natural-order unpivoted LDL' + substitution (the tuned kernels: qp_tpp.cuh, qp_wdm.cuh, qp_warp.cuh)
 *   3  as 2 followed by one step of iterative refinement against the assembled KKT matrix */
static int g_solve_mode = 0;
void oracle_set_solve_mode(int mode) { g_solve_mode = mode; }

static int g_diag_n = 0; /* H-block size for the ordering modes */
static int g_variant = 0; /* diagnostic arithmetic variants: 1 = no FMA, 2 = divide by d instead of multiplying by 1/d,
                             4 = left-looking (dot-product) accumulation order, 8 = solve divides by d */
void oracle_set_variant(int v) { g_variant = v; }
static double vfma(double a, double b, double c) { return (g_variant & 1) ? (a * b + c) : fma(a, b, c); }
static void unpivoted_ldlt_solve(const double* H, int p, double* x_io, double* scratch) {
  /* scratch: p*p; lower triangle of H significant. Modes >= 5 factor P K P' with a STATIC permutation chosen from the
   * assembled diagonal: 5 = H block sorted by descending diagonal (= Eigen's order: its pivot search sees the
   * not-yet-updated diagonal, LDLT.h:324-326, so equality rows (diagonal 0) come last in natural order);
   * 6 = variables with diagonal > 1e6 first, natural order otherwise; 7 = natural order, but the LAST TWO pivots are
   * taken larger-|diagonal|-first dynamically. */
  double* L = scratch;
  int perm[256];
  double xp[256];
  const int n = g_diag_n;
  for (int i = 0; i < p; ++i) perm[i] = i;
  if (g_solve_mode == 5) {
    for (int i = 1; i < n; ++i) { /* stable insertion sort, descending diagonal */
      const int v = perm[i];
      const double key = H[(size_t)v * p + v];
      int j = i - 1;
      while (j >= 0 && H[(size_t)perm[j] * p + perm[j]] < key) { perm[j + 1] = perm[j]; --j; }
      perm[j + 1] = v;
    }
  } else if (g_solve_mode == 6) {
    int cnt = 0, tmp[256];
    for (int i = 0; i < n; ++i) if (H[(size_t)i * p + i] > 1e6) tmp[cnt++] = i;
    for (int i = 0; i < n; ++i) if (!(H[(size_t)i * p + i] > 1e6)) tmp[cnt++] = i;
    for (int i = 0; i < n; ++i) perm[i] = tmp[i];
  }
  for (int j = 0; j < p; ++j)
    for (int i = j; i < p; ++i) {
      const int a = perm[i], b = perm[j];
      L[(size_t)j * p + i] = a >= b ? H[(size_t)b * p + a] : H[(size_t)a * p + b];
    }
  for (int i = 0; i < p; ++i) xp[i] = x_io[perm[i]];
  double* x = xp;
  int swapped = 0;
  for (int j = 0; j < p; ++j) {
    double t[256];
    if (g_solve_mode == 7 && j == p - 2 && fabs(L[(size_t)(p - 1) * p + (p - 1)]) > fabs(L[(size_t)j * p + j])) {
      /* swap the last two rows/columns */
      for (int c = 0; c < j; ++c) { double v = L[(size_t)c * p + j]; L[(size_t)c * p + j] = L[(size_t)c * p + p - 1]; L[(size_t)c * p + p - 1] = v; }
      { double v = L[(size_t)j * p + j]; L[(size_t)j * p + j] = L[(size_t)(p - 1) * p + p - 1]; L[(size_t)(p - 1) * p + p - 1] = v; }
      { double v = x[j]; x[j] = x[p - 1]; x[p - 1] = v; }
      swapped = 1;
    }
    if (g_variant & 4) { /* left-looking: apply all earlier columns to column j now (Eigen's bordered form, LDLT.h:356-362) */
      double tmp[256], dot = 0.0;
      for (int c = 0; c < j; ++c) { tmp[c] = (1.0 / L[(size_t)c * p + c]) * L[(size_t)c * p + j]; }
      for (int c = 0; c < j; ++c) { tmp[c] = L[(size_t)c * p + j] / L[(size_t)c * p + c]; } /* D_c * L_jc, D stored inverted */
      for (int c = 0; c < j; ++c) dot = vfma(L[(size_t)c * p + j], tmp[c], dot);
      L[(size_t)j * p + j] -= dot;
      for (int i = j + 1; i < p; ++i) {
        double acc = 0.0;
        for (int c = 0; c < j; ++c) acc = vfma(L[(size_t)c * p + i], tmp[c], acc);
        L[(size_t)j * p + i] -= acc;
      }
      const double d = L[(size_t)j * p + j];
      for (int i = j + 1; i < p; ++i) L[(size_t)j * p + i] = (g_variant & 2) ? L[(size_t)j * p + i] / d : L[(size_t)j * p + i] * (1.0 / d);
      L[(size_t)j * p + j] = 1.0 / d;
      continue;
    }
    const double dj = L[(size_t)j * p + j];
    const double inv = 1.0 / dj;
    for (int i = j + 1; i < p; ++i) t[i] = L[(size_t)j * p + i];
    for (int i = j + 1; i < p; ++i) {
      const double l = (g_variant & 2) ? t[i] / dj : t[i] * inv;
      for (int c = j + 1; c <= i; ++c) L[(size_t)c * p + i] = vfma(-l, t[c], L[(size_t)c * p + i]);
      L[(size_t)j * p + i] = l;
    }
    L[(size_t)j * p + j] = inv;
  }
  for (int j = 0; j < p; ++j)
    for (int i = j + 1; i < p; ++i) x[i] = vfma(-L[(size_t)j * p + i], x[j], x[i]);
  for (int j = 0; j < p; ++j) x[j] = (g_variant & 8) ? x[j] / (1.0 / L[(size_t)j * p + j]) : x[j] * L[(size_t)j * p + j];
  for (int j = p - 1; j >= 0; --j)
    for (int i = j + 1; i < p; ++i) x[j] = vfma(-L[(size_t)j * p + i], x[i], x[j]);
  if (swapped) { double v = x[p - 2]; x[p - 2] = x[p - 1]; x[p - 1] = v; }
  for (int i = 0; i < p; ++i) x_io[perm[i]] = x[i];
}

/* The four block mat-vecs against H^-1 shared by SolveForUpdate (qp.cc:350-356) and
 * SolveForUpdateNoInequalities (qp.cc:379-385). */
static void apply_inverse(const oq_problem* q, oq_work* w, const double* r_x, const double* r_pe) {
  const int n = q->n, k = q->k, p = w->p;
  double *dx = XB(w->delta, q), *dy = YB(w->delta, q);
  if (g_solve_mode != 0) { /* [dx; -dy] = K^-1 [-r_x; -r_pe] by substitution */
    double u[256], rhs[256];
    for (int i = 0; i < n; ++i) u[i] = -r_x[i];
    for (int i = 0; i < k; ++i) u[n + i] = -r_pe[i];
    memcpy(rhs, u, sizeof(double) * p);
    g_diag_n = n;
    if (g_solve_mode == 1) {
      eigen_ldlt_solve_inplace(w->LD, w->transp, p, u, 1);
    } else {
      unpivoted_ldlt_solve(w->H, p, u, w->Hinv);
      if (g_solve_mode == 3) {
        double res[256];
        for (int i = 0; i < p; ++i) {
          double acc = rhs[i];
          for (int j = 0; j < p; ++j) {
            const double h = i >= j ? w->H[(size_t)j * p + i] : w->H[(size_t)i * p + j];
            acc = fma(-h, u[j], acc);
          }
          res[i] = acc;
        }
        unpivoted_ldlt_solve(w->H, p, res, w->Hinv);
        for (int i = 0; i < p; ++i) u[i] += res[i];
      }
    }
    for (int i = 0; i < n; ++i) dx[i] = u[i];
    for (int i = 0; i < k; ++i) dy[i] = -u[n + i];
    return;
  }
  for (int i = 0; i < n; ++i) {
    double acc = 0.0;
    for (int j = 0; j < n; ++j) acc += w->Hinv[(size_t)j * p + i] * -r_x[j];
    dx[i] = acc;
  }
  if (k > 0) {
    for (int i = 0; i < n; ++i) {
      double acc = 0.0;
      for (int j = 0; j < k; ++j) acc += w->Hinv[(size_t)(n + j) * p + i] * -r_pe[j];
      dx[i] += acc;
    }
    for (int i = 0; i < k; ++i) {
      double acc = 0.0;
      for (int j = 0; j < n; ++j) acc += w->Hinv[(size_t)j * p + n + i] * r_x[j];
      dy[i] = acc;
      acc = 0.0;
      for (int j = 0; j < k; ++j) acc += w->Hinv[(size_t)(n + j) * p + n + i] * r_pe[j];
      dy[i] += acc;
    }
  }
}

/* SolveForUpdate, qp.cc:318-364 */
static void solve_for_update(const oq_problem* q, oq_work* w, double mu) {
  const int n = q->n, m = q->m;
  const double *s = SB(w->vars, q), *z = ZB(w->vars, q);
  const double *r_d = XB(w->r, q), *r_comp = SB(w->r, q), *r_pe = YB(w->r, q), *r_pi = ZB(w->r, q);
  const double *ds_aff = SB(w->delta_aff, q), *dz_aff = ZB(w->delta_aff, q);
  memcpy(w->r_aug, r_d, sizeof(double) * n);
  for (int i = 0; i < m; ++i) {
    const int v = con_var(q, i);
    const double a = con_a(q, i);
    w->r_aug[v] += a * (z[i] / s[i]) * r_pi[i];
    w->r_aug[v] += a * (r_comp[i] + (ds_aff[i] * dz_aff[i]) - mu) / s[i];
  }
  apply_inverse(q, w, w->r_aug, r_pe);
  const double* dx = XB(w->delta, q);
  double *ds = SB(w->delta, q), *dz = ZB(w->delta, q);
  for (int i = 0; i < m; ++i) {
    const int v = con_var(q, i);
    const double a = con_a(q, i);
    ds[i] = a * dx[v] + r_pi[i];
    dz[i] = -(z[i] / s[i]) * ds[i] - (1 / s[i]) * (r_comp[i] + (ds_aff[i] * dz_aff[i]) - mu);
  }
}

/* ComputeAlpha, qp.cc:491-507 */
static double compute_alpha(const double* val, const double* dval, int len, double tau) {
  double alpha = 1.0;
  for (int i = 0; i < len; ++i) {
    const double updated = val[i] + dval[i];
    if (updated <= 0.0 && fabs(dval[i]) > 0) {
      const double cand = -tau * val[i] / dval[i];
      if (cand < alpha) alpha = cand;
    }
  }
  return alpha;
}

/* ComputeMu, qp.cc:509-516 */
static double compute_mu(const oq_problem* q, const oq_work* w) {
  if (q->m == 0) return 0.0;
  const double *s = SB(w->vars, q), *z = ZB(w->vars, q);
  double dot = 0.0;
  for (int i = 0; i < q->m; ++i) dot += s[i] * z[i];
  return dot / (double)q->m;
}

static double norm2(const double* v, int len) {
  double acc = 0.0;
  for (int i = 0; i < len; ++i) acc += v[i] * v[i];
  return sqrt(acc);
}

/* ComputeErrors, qp.cc:423-437; out = {r_dual, r_comp, r_primal_eq, r_primal_ineq} */
static void compute_errors(const oq_problem* q, const oq_work* w, double mu, double* out) {
  out[0] = norm2(XB(w->r, q), q->n);
  out[1] = 0.0; out[2] = 0.0; out[3] = 0.0;
  if (q->k > 0) out[2] = norm2(YB(w->r, q), q->k);
  if (q->m > 0) {
    const double* rc = SB(w->r, q);
    double sq = 0.0, sum = 0.0;
    for (int i = 0; i < q->m; ++i) { sq += rc[i] * rc[i]; sum += rc[i]; }
    const double corrected = sq - 2 * (sum * mu) + (mu * mu) * q->m;
    out[1] = sqrt(corrected > 0. ? corrected : 0.);
    out[3] = norm2(ZB(w->r, q), q->m);
  }
}

static double max4(const double* e) {
  double v = e[0];
  for (int i = 1; i < 4; ++i) if (e[i] > v) v = e[i];
  return v;
}

/* LinearInequalityConstraint::ClampX, qp.hpp:43-53 */
static double clamp_x(double a, double b, double x) {
  if (a < 0) { const double hi = b / -a; return x < hi ? x : hi; }
  const double lo = -b / a;
  return x > lo ? x : lo;
}

/* ComputeInitialGuess, qp.cc:439-482. Returns 0 or a failure status. */
static int initial_guess(const oq_problem* q, oq_work* w, int method) {
  const int n = q->n, k = q->k, m = q->m;
  if (method == MO_GUESS_USER_PROVIDED) return 0;
  double *x = XB(w->vars, q), *s = SB(w->vars, q), *y = YB(w->vars, q), *z = ZB(w->vars, q);
  memset(x, 0, sizeof(double) * n);
  memset(y, 0, sizeof(double) * k);
  if (method == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED) {
    const int st = compute_ldlt(q, w, 0);
    if (st) return st;
    evaluate_kkt(q, w, 0);
    apply_inverse(q, w, XB(w->r, q), YB(w->r, q)); /* SolveForUpdateNoInequalities, qp.cc:366-386 */
    memcpy(x, XB(w->delta, q), sizeof(double) * n);
    memcpy(y, YB(w->delta, q), sizeof(double) * k);
  }
  for (int i = 0; i < m; ++i) { /* sequential clamp in list order, qp.cc:464-467 */
    const int v = con_var(q, i);
    x[v] = clamp_x(con_a(q, i), q->ib[i], x[v]);
  }
  for (int i = 0; i < m; ++i) { /* qp.cc:472-481 */
    const double sv = con_a(q, i) * x[con_var(q, i)] + q->ib[i];
    s[i] = 1.0e-9 > sv ? 1.0e-9 : sv;
    z[i] = 1.0 / s[i];
  }
  return 0;
}

/* Iterate, qp.cc:153-201. ip = {mu, alpha.primal, alpha.dual, probe.primal, probe.dual, mu_affine}. */
static int iterate(const oq_problem* q, oq_work* w, double mu_input, int strategy, double* ip) {
  const int n = q->n, k = q->k, m = q->m;
  evaluate_kkt(q, w, 1);
  ip[0] = mu_input; ip[1] = 1.0; ip[2] = 1.0; ip[3] = NAN; ip[4] = NAN; ip[5] = NAN;
  const int st = compute_ldlt(q, w, 1);
  if (st) return st;
  double *s = SB(w->vars, q), *z = ZB(w->vars, q);
  if (m == 0) {
    solve_for_update(q, w, 0.0);
  } else if (strategy != MO_BARRIER_PREDICTOR_CORRECTOR) {
    solve_for_update(q, w, ip[0]);
  } else {
    solve_for_update(q, w, 0.0);
    ip[3] = compute_alpha(s, SB(w->delta, q), m, 1.0);
    ip[4] = compute_alpha(z, ZB(w->delta, q), m, 1.0);
    memcpy(w->delta_aff, w->delta, sizeof(double) * (n + 2 * m + k));
    /* ComputePredictorCorrectorMuAffine, qp.cc:519-537 */
    const double *ds = SB(w->delta_aff, q), *dz = ZB(w->delta_aff, q);
    double sdz = 0, zds = 0, dsdz = 0;
    for (int i = 0; i < m; ++i) { sdz += s[i] * dz[i]; zds += z[i] * ds[i]; dsdz += ds[i] * dz[i]; }
    double mu_aff = ip[0];
    mu_aff += ip[4] * sdz / (double)m;
    mu_aff += ip[3] * zds / (double)m;
    mu_aff += (ip[4] * ip[3]) * dsdz / (double)m;
    ip[5] = mu_aff > 0.0 ? mu_aff : 0.0;
    const double sigma = pow(ip[5] / ip[0], 3);
    ip[0] = sigma * mu_input;
    solve_for_update(q, w, ip[0]);
  }
  if (m > 0) {
    ip[1] = compute_alpha(s, SB(w->delta, q), m, 0.995);
    ip[2] = compute_alpha(z, ZB(w->delta, q), m, 0.995);
  }
  double *x = XB(w->vars, q), *y = YB(w->vars, q);
  for (int i = 0; i < n; ++i) x[i] += XB(w->delta, q)[i] * ip[1];
  for (int i = 0; i < m; ++i) s[i] += SB(w->delta, q)[i] * ip[1];
  for (int i = 0; i < k; ++i) y[i] += YB(w->delta, q)[i] * ip[2];
  for (int i = 0; i < m; ++i) z[i] += ZB(w->delta, q)[i] * ip[2];
  return 0;
}

static void bind_problem(const mo_qp_batch* b, int64_t p, oq_problem* q) {
  q->n = b->n; q->k = b->k;
  q->m = b->ineq_count ? b->ineq_count[p] : b->m;
  q->G = b->G + p * b->n * b->n;
  q->c = b->c + p * b->n;
  q->A = b->k ? b->A_eq + p * b->k * b->n : NULL;
  q->b = b->k ? b->b_eq + p * b->k : NULL;
  q->box_pairs = (b->flags & MO_QP_FLAG_BOX_PAIRS) != 0;
  q->var = (b->ineq_var && !q->box_pairs) ? b->ineq_var + p * b->m : NULL;
  q->ia = (b->ineq_a && !q->box_pairs) ? b->ineq_a + p * b->m : NULL;
  q->ib = b->m ? b->ineq_b + p * b->m : NULL;
}

/* compact [x|s(mc)|y|z(mc)] <-> fixed stride [x|s(m)|y|z(m)] (unused slots NaN) */
static void store_vars(const mo_qp_batch* b, const oq_problem* q, const double* v, double* dst) {
  const int n = b->n, k = b->k, m = b->m, mc = q->m;
  for (int i = 0; i < n; ++i) dst[i] = v[i];
  for (int i = 0; i < m; ++i) dst[n + i] = i < mc ? v[n + i] : NAN;
  for (int i = 0; i < k; ++i) dst[n + m + i] = v[n + mc + i];
  for (int i = 0; i < m; ++i) dst[n + m + k + i] = i < mc ? v[n + mc + k + i] : NAN;
}

static void load_vars(const mo_qp_batch* b, const oq_problem* q, const double* src, double* v) {
  const int n = b->n, k = b->k, m = b->m, mc = q->m;
  for (int i = 0; i < n; ++i) v[i] = src[i];
  for (int i = 0; i < mc; ++i) v[n + i] = src[n + i];
  for (int i = 0; i < k; ++i) v[n + mc + i] = src[n + m + i];
  for (int i = 0; i < mc; ++i) v[n + mc + k + i] = src[n + m + k + i];
}

/* Solve, qp.cc:100-151, for every problem of a HOST batch. Returns total iterations or <0. */
int64_t oracle_qp_solve_batch(const mo_qp_batch* b, const mo_qp_params* prm, const mo_qp_out* out) {
  if (!b || !prm || !out || !out->vars || !out->status || !out->iters) return MO_ERR_INVALID_ARGUMENT;
  /* CheckParams, qp.cc:76-82 */
  if (!(prm->initial_mu > 0) || !(prm->sigma > 0) || !(prm->sigma <= 1.0) || !(prm->termination_kkt_tol > 0) ||
      !(prm->max_iterations > 0))
    return MO_ERR_INVALID_PARAMS;
  const int64_t stride = b->n + 2 * b->m + b->k;
  int64_t total = 0;
  for (int64_t p = 0; p < b->batch; ++p) {
    oq_problem q;
    bind_problem(b, p, &q);
    for (int i = 0; i < q.m; ++i)
      if (con_var(&q, i) < 0 || con_var(&q, i) >= q.n) return MO_ERR_INVALID_ARGUMENT; /* qp.cc:70-72 */
    oq_work w;
    work_alloc(&w, &q);
    if (prm->initial_guess_method == MO_GUESS_USER_PROVIDED) load_vars(b, &q, out->vars + p * stride, w.vars);
    int status = initial_guess(&q, &w, prm->initial_guess_method);
    int iters = 0;
    double kkt_prev[4] = {0, 0, 0, 0}, kkt_after[4] = {0, 0, 0, 0}, ip[6];
    if (!status) {
      status = MO_QP_MAX_ITERATIONS;
      evaluate_kkt(&q, &w, 1);
      double mu = prm->initialize_mu_with_complementarity ? compute_mu(&q, &w) : prm->initial_mu;
      for (int iter = 0; iter < prm->max_iterations; ++iter) {
        compute_errors(&q, &w, mu, kkt_prev);
        const int st = iterate(&q, &w, mu, prm->barrier_strategy, ip);
        if (st) { status = st; break; }
        evaluate_kkt(&q, &w, 1);
        compute_errors(&q, &w, mu, kkt_after);
        if (out->iter_log) {
          double* row = out->iter_log + (p * prm->max_iterations + iter) * MO_QP_ITER_LOG_STRIDE;
          memcpy(row, kkt_prev, sizeof(kkt_prev));
          memcpy(row + 4, kkt_after, sizeof(kkt_after));
          memcpy(row + 8, ip, sizeof(ip));
        }
        ++iters;
        if (max4(kkt_after) < prm->termination_kkt_tol && compute_mu(&q, &w) < prm->termination_complementarity_tol) {
          status = MO_QP_SATISFIED_KKT_TOL;
          break;
        }
        if (max4(kkt_after) <= mu || !prm->decrease_mu_only_on_small_error) {
          if (prm->barrier_strategy == MO_BARRIER_FIXED_DECREASE) mu *= prm->sigma;
          else mu = prm->sigma * compute_mu(&q, &w);
        }
      }
    }
    store_vars(b, &q, w.vars, out->vars + p * stride);
    out->status[p] = status;
    out->iters[p] = iters;
    if (out->kkt_final) memcpy(out->kkt_final + p * 4, kkt_after, sizeof(kkt_after));
    if (out->lagrange) { /* ComputeLagrangeMultiplierSummary, qp.cc:539-546 */
      double mn = NAN, li = NAN;
      if (q.k > 0) {
        const double* y = YB(w.vars, &q);
        mn = y[0]; li = fabs(y[0]);
        for (int i = 1; i < q.k; ++i) { if (y[i] < mn) mn = y[i]; if (fabs(y[i]) > li) li = fabs(y[i]); }
      }
      out->lagrange[p * 2] = mn;
      out->lagrange[p * 2 + 1] = li;
    }
    total += iters;
    work_free(&w);
  }
  return total;
}

/* One Iterate() per problem on caller state (same contract as mo_qp_ip_newton_step_batch). */
int oracle_qp_newton_step_batch(const mo_qp_batch* b, double mu, int32_t strategy, double* vars, double* delta_out,
                                double* ip_out, double* kkt_out, int32_t* status) {
  if (!b || !vars || !status) return MO_ERR_INVALID_ARGUMENT;
  const int64_t stride = b->n + 2 * b->m + b->k;
  for (int64_t p = 0; p < b->batch; ++p) {
    oq_problem q;
    bind_problem(b, p, &q);
    oq_work w;
    work_alloc(&w, &q);
    load_vars(b, &q, vars + p * stride, w.vars);
    double ip[6];
    const int st = iterate(&q, &w, mu, strategy, ip);
    status[p] = st;
    if (!st) {
      store_vars(b, &q, w.vars, vars + p * stride);
      if (delta_out) store_vars(b, &q, w.delta, delta_out + p * stride);
      if (ip_out) memcpy(ip_out + p * 6, ip, sizeof(ip));
      if (kkt_out) {
        evaluate_kkt(&q, &w, 1);
        compute_errors(&q, &w, mu, kkt_out + p * 4);
      }
    }
    work_free(&w);
  }
  return MO_OK;
}
