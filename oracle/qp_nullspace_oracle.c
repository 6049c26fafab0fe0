/*
 * TEST INFRASTRUCTURE — not part of the product (see oracle/qp_ip_oracle.c).
 *
 * Plain-C restatement of QPNullSpaceSolver::Solve (gareth-cross/mini_opt, source/qp.cc:679-729) INCLUDING the parts of
 * Eigen 3.4.0 it leans on: ColPivHouseholderQR<MatrixXd>::computeInPlace (Eigen/src/QR/ColPivHouseholderQR.h:480-580:
 * column pivoting on the running column norms with the LAPACK-style downdate, makeHouseholderInPlace of
 * Eigen/src/Householder/Householder.h:67-96), rank() (:398-407), householderQ() evaluated densely, the triangular solve
 * of R1^T, and LLT<Lower> (Eigen/src/Cholesky/LLT.h:290-316).
 *
 *   [Q1 Q2] R = A_eq^T P,  u = Q1 R1^-T P^T (-b_eq),  (Q2^T G Q2) y = -Q2^T (c + G u),  x = u + Q2 y
 *
 * status: 0 SUCCESS, 1 NOT_POSITIVE_DEFINITE (QPNullSpaceTerminationState, structs.hpp:137-142),
 *         2 rank(A_eq) < k: the reference multiplies mismatched blocks there (an Eigen assertion, compiled out): batched
 *           stand-in, x = 0.
 * Parity: PINNED against oracle/_ref (ref_qp_nullspace_batch) in tests/test_oracle_qp.py.
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "mini_opt_b200.h"

static double gsym(const double* G, int n, int i, int j) { return i >= j ? G[(size_t)j * n + i] : G[(size_t)i * n + j]; }

/* One problem. G n x n column-major (lower triangle read), A k x n column-major. */
int oracle_qp_nullspace_one(int n, int k, const double* G, const double* c, const double* A, const double* b, double* x) {
  const int rows = n, cols = k, size = n < k ? n : k;
  double* qr = (double*)calloc((size_t)rows * cols + 1, sizeof(double)); /* A^T, rows x cols column-major */
  double* Q = (double*)calloc((size_t)rows * rows + 1, sizeof(double));
  double* hc = (double*)calloc((size_t)size + 1, sizeof(double));
  double* upd = (double*)calloc((size_t)cols + 1, sizeof(double));
  double* dir = (double*)calloc((size_t)cols + 1, sizeof(double));
  double* tmp = (double*)calloc((size_t)rows + cols + 1, sizeof(double));
  int* transp = (int*)calloc((size_t)cols + 1, sizeof(int));
  int* perm = (int*)calloc((size_t)cols + 1, sizeof(int));
  int status = 0;
  for (int j = 0; j < cols; ++j)
    for (int i = 0; i < rows; ++i) qr[(size_t)j * rows + i] = A[(size_t)i * k + j]; /* A^T(i, j) = A(j, i) */
#define QR(i, j) qr[(size_t)(j) * rows + (i)]
#define QM(i, j) Q[(size_t)(j) * rows + (i)]
  double maxnorm = 0.0;
  for (int j = 0; j < cols; ++j) {
    double s = 0.0;
    for (int i = 0; i < rows; ++i) s += QR(i, j) * QR(i, j);
    upd[j] = dir[j] = sqrt(s);
    if (upd[j] > maxnorm) maxnorm = upd[j];
  }
  const double th = maxnorm * DBL_EPSILON / (double)rows, threshold_helper = th * th;
  const double downdate = sqrt(DBL_EPSILON);
  int nonzero_pivots = size;
  double maxpivot = 0.0;
  for (int kk = 0; kk < size; ++kk) {
    int big = kk;
    double best = upd[kk];
    for (int j = kk + 1; j < cols; ++j)
      if (upd[j] > best) { best = upd[j]; big = j; }
    if (nonzero_pivots == size && best * best < threshold_helper * (double)(rows - kk)) nonzero_pivots = kk;
    transp[kk] = big;
    if (big != kk) {
      for (int i = 0; i < rows; ++i) { double t = QR(i, kk); QR(i, kk) = QR(i, big); QR(i, big) = t; }
      { double t = upd[kk]; upd[kk] = upd[big]; upd[big] = t; }
      { double t = dir[kk]; dir[kk] = dir[big]; dir[big] = t; }
    }
    /* makeHouseholderInPlace on QR(kk.., kk) */
    double tail = 0.0;
    for (int i = kk + 1; i < rows; ++i) tail += QR(i, kk) * QR(i, kk);
    const double c0 = QR(kk, kk);
    double beta, tau;
    if (tail <= DBL_MIN) {
      tau = 0.0; beta = c0;
      for (int i = kk + 1; i < rows; ++i) QR(i, kk) = 0.0;
    } else {
      beta = sqrt(c0 * c0 + tail);
      if (c0 >= 0.0) beta = -beta;
      for (int i = kk + 1; i < rows; ++i) QR(i, kk) /= (c0 - beta);
      tau = (beta - c0) / beta;
    }
    hc[kk] = tau;
    QR(kk, kk) = beta;
    if (fabs(beta) > maxpivot) maxpivot = fabs(beta);
    /* apply H = I - tau v v^T (v = [1; essential]) to the trailing columns: applyHouseholderOnTheLeft */
    for (int j = kk + 1; j < cols; ++j) {
      double t = 0.0;
      for (int i = kk + 1; i < rows; ++i) t += QR(i, kk) * QR(i, j);
      t += QR(kk, j);
      QR(kk, j) -= tau * t;
      for (int i = kk + 1; i < rows; ++i) QR(i, j) -= (tau * QR(i, kk)) * t;
    }
    for (int j = kk + 1; j < cols; ++j) { /* norm downdate, ColPivHouseholderQR.h:551-571 */
      if (upd[j] != 0.0) {
        double t = fabs(QR(kk, j)) / upd[j];
        t = (1.0 + t) * (1.0 - t);
        t = t < 0.0 ? 0.0 : t;
        const double ratio = upd[j] / dir[j], t2 = t * ratio * ratio;
        if (t2 <= downdate) {
          double s = 0.0;
          for (int i = kk + 1; i < rows; ++i) s += QR(i, j) * QR(i, j);
          dir[j] = sqrt(s);
          upd[j] = dir[j];
        } else {
          upd[j] *= sqrt(t);
        }
      }
    }
  }
  /* rank(), ColPivHouseholderQR.h:398-407 */
  int rank = 0;
  {
    const double pre = fabs(maxpivot) * (DBL_EPSILON * (double)size);
    for (int i = 0; i < nonzero_pivots; ++i) rank += fabs(QR(i, i)) > pre;
  }
  for (int i = 0; i < n; ++i) x[i] = 0.0;
  if (rank < k) { status = 2; goto done; }
  /* permutation indices: identity with the transpositions applied on the right in order */
  for (int j = 0; j < cols; ++j) perm[j] = j;
  for (int kk = 0; kk < size; ++kk) { int t = perm[kk]; perm[kk] = perm[transp[kk]]; perm[transp[kk]] = t; }
  /* Q = H_0 ... H_{size-1} applied to the identity (HouseholderSequence::evalTo) */
  for (int i = 0; i < rows; ++i) QM(i, i) = 1.0;
  for (int kk = size - 1; kk >= 0; --kk) {
    for (int j = kk; j < rows; ++j) {
      double t = QM(kk, j);
      for (int i = kk + 1; i < rows; ++i) t += QR(i, kk) * QM(i, j);
      QM(kk, j) -= hc[kk] * t;
      for (int i = kk + 1; i < rows; ++i) QM(i, j) -= (hc[kk] * QR(i, kk)) * t;
    }
  }
  {
    /* u = Q1 R1^-T P^T (-b): forward substitution with the lower triangular R1^T */
    double* w = tmp;        /* k */
    double* u = tmp + cols; /* n */
    for (int i = 0; i < k; ++i) w[i] = -b[perm[i]];
    for (int i = 0; i < k; ++i) {
      double acc = w[i];
      for (int j = 0; j < i; ++j) acc -= QR(j, i) * w[j]; /* R^T(i, j) = R(j, i) */
      w[i] = acc / QR(i, i);
    }
    for (int i = 0; i < n; ++i) {
      double acc = 0.0;
      for (int j = 0; j < k; ++j) acc += QM(i, j) * w[j];
      u[i] = acc;
    }
    const int nz = n - k;
    if (nz == 0) {
      for (int i = 0; i < n; ++i) x[i] = u[i];
      goto done;
    }
    double* GZ = (double*)calloc((size_t)n * nz + 1, sizeof(double));
    double* Gr = (double*)calloc((size_t)nz * nz + 1, sizeof(double));
    double* y = (double*)calloc((size_t)nz + 1, sizeof(double));
    double* gu = (double*)calloc((size_t)n + 1, sizeof(double));
    for (int j = 0; j < nz; ++j)
      for (int i = 0; i < n; ++i) {
        double acc = 0.0;
        for (int l = 0; l < n; ++l) acc += gsym(G, n, i, l) * QM(l, k + j);
        GZ[(size_t)j * n + i] = acc;
      }
    for (int j = 0; j < nz; ++j)
      for (int i = 0; i < nz; ++i) {
        double acc = 0.0;
        for (int l = 0; l < n; ++l) acc += QM(l, k + i) * GZ[(size_t)j * n + l];
        Gr[(size_t)j * nz + i] = acc;
      }
    /* LLT<Lower> unblocked, LLT.h:290-316 */
    for (int kk = 0; kk < nz && status == 0; ++kk) {
      double d = Gr[(size_t)kk * nz + kk];
      for (int j = 0; j < kk; ++j) d -= Gr[(size_t)j * nz + kk] * Gr[(size_t)j * nz + kk];
      if (d <= 0.0) { status = 1; break; }
      d = sqrt(d);
      Gr[(size_t)kk * nz + kk] = d;
      for (int i = kk + 1; i < nz; ++i) {
        double acc = Gr[(size_t)kk * nz + i];
        for (int j = 0; j < kk; ++j) acc -= Gr[(size_t)j * nz + i] * Gr[(size_t)j * nz + kk];
        Gr[(size_t)kk * nz + i] = acc / d;
      }
    }
    if (status == 0) {
      for (int i = 0; i < n; ++i) {
        double acc = c[i];
        for (int l = 0; l < n; ++l) acc += gsym(G, n, i, l) * u[l];
        gu[i] = acc;
      }
      for (int j = 0; j < nz; ++j) {
        double acc = 0.0;
        for (int l = 0; l < n; ++l) acc += QM(l, k + j) * gu[l];
        y[j] = -acc;
      }
      for (int i = 0; i < nz; ++i) { /* L y' = y */
        double acc = y[i];
        for (int j = 0; j < i; ++j) acc -= Gr[(size_t)j * nz + i] * y[j];
        y[i] = acc / Gr[(size_t)i * nz + i];
      }
      for (int i = nz - 1; i >= 0; --i) { /* L^T y = y' */
        double acc = y[i];
        for (int j = i + 1; j < nz; ++j) acc -= Gr[(size_t)i * nz + j] * y[j];
        y[i] = acc / Gr[(size_t)i * nz + i];
      }
      for (int i = 0; i < n; ++i) {
        double acc = u[i];
        for (int j = 0; j < nz; ++j) acc += QM(i, k + j) * y[j];
        x[i] = acc;
      }
    }
    free(GZ); free(Gr); free(y); free(gu);
  }
done:
#undef QR
#undef QM
  free(qr); free(Q); free(hc); free(upd); free(dir); free(tmp); free(transp); free(perm);
  return status;
}

/* x [batch][n], status [batch]; reads G, c, A_eq, b_eq of the batch (inequalities ignored). */
int oracle_qp_nullspace_batch(const mo_qp_batch* B, double* x, int32_t* status) {
  if (!B || !x || !status || B->k <= 0) return MO_ERR_INVALID_ARGUMENT;
  const int n = B->n, k = B->k;
  for (int64_t p = 0; p < B->batch; ++p)
    status[p] = oracle_qp_nullspace_one(n, k, B->G + p * n * n, B->c + p * n, B->A_eq + p * k * n, B->b_eq + p * k, x + p * n);
  return MO_OK;
}
