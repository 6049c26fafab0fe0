// Generic interior point QP kernel: one problem per thread block (THREADS = 32, 128 or 256), runtime
// dimensions, general (variable, a, b) inequality lists with per-problem counts, every barrier strategy
// and initial-guess method of the reference, optional per-iteration log.
//
// This is the coverage / audit path behind the same C ABI as the tuned kernels: the reduced KKT matrix
// lives in shared memory and is factored with the SAME pivot rule as Eigen's LDLT<MatrixXd, Lower>
// (largest |diagonal| of the trailing block, first wins; dependencies/eigen/Eigen/src/Cholesky/LDLT.h:297-403)
// that QPInteriorPointSolver::ComputeLDLT uses (source/qp.cc:302), so it accepts exactly the problems
// the reference accepts (semi-definite G made non-singular by the equality block, zero pivots, ...).
// Unlike the reference it solves for the right-hand sides it needs instead of forming H^-1 (qp.cc:310).
#pragma once

#include <float.h>

#include "qp_common.cuh"

namespace mo {

// Index of this thread inside its problem's team: the lane for THREADS == 32 (several one-warp teams may share a CTA,
// e.g. the lock-step IK kernel), threadIdx.x for the one-CTA-per-problem instantiations.
template <int THREADS>
__device__ __forceinline__ int team_tid() { return THREADS == 32 ? (int)(threadIdx.x & 31) : (int)threadIdx.x; }

template <int THREADS>
__device__ __forceinline__ void block_sync() {
  if (THREADS == 32) __syncwarp(); else __syncthreads();
}

// Sum over the block, same value returned to every thread. `scratch` holds >= THREADS/32 doubles.
template <int THREADS>
__device__ __forceinline__ double block_sum(double v, double* scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (THREADS == 32) return v;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int w = 0; w < THREADS / 32; ++w) t += scratch[w];
  return t;
}

template <int THREADS>
__device__ __forceinline__ double block_min(double v, double* scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (THREADS == 32) return v;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  double t = scratch[0];
#pragma unroll
  for (int w = 1; w < THREADS / 32; ++w) t = fmin(t, scratch[w]);
  return t;
}

// argmax of (key, smallest index wins ties); returns the index to every thread.
template <int THREADS>
__device__ __forceinline__ int block_argmax(double key, int idx, double* scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ok = __shfl_xor_sync(0xffffffffu, key, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ok > key || (ok == key && oi < idx)) { key = ok; idx = oi; }
  }
  if (THREADS == 32) return idx;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int* iscratch = reinterpret_cast<int*>(scratch + THREADS / 32);
  __syncthreads();
  if (lane == 0) { scratch[warp] = key; iscratch[warp] = idx; }
  __syncthreads();
  double bk = scratch[0];
  int bi = iscratch[0];
#pragma unroll
  for (int w = 1; w < THREADS / 32; ++w) {
    const double ok = scratch[w];
    const int oi = iscratch[w];
    if (ok > bk || (ok == bk && oi < bi)) { bk = ok; bi = oi; }
  }
  return bi;
}

// Shared-memory workspace of one problem.
struct GenWork {
  int n, k, m, p, ld, T;  // m = constraints of THIS problem; T = n + 2m + k (compact layout)
  double* Kmat;           // p x ld, column-major, lower triangle significant
  double* vars;           // [x | s | y | z]  (qp.cc:548-582)
  double* r;              // [r_d | r_comp | r_pe | r_pi]
  double* delta;          // [dx | ds | dy | dz]
  double* daff;           // predictor step (PREDICTOR_CORRECTOR only)
  double* rhs;            // p
  double* temp;           // p   (LDLT m_temporary)
  double* w;              // m   z/s
  double* ca;             // m   constraint a
  double* cb;             // m   constraint b
  double* scratch;        // 2 * THREADS/32 + 2
  int* cv;                // m   constraint variable
  int* transp;            // p
};


#define KM(i, j) W.Kmat[(size_t)(j) * W.ld + (i)]

// Diagonal-pivoted LDL' of the lower triangle of W.Kmat, in place (same elimination order and pivot rule
// as Eigen's ldlt_inplace<Lower>::unblocked). Returns true on success (Eigen's `ret`), uniformly.
template <int THREADS>
__device__ bool gen_ldlt_factor(const GenWork& W) {
  const int tid = team_tid<THREADS>(), size = W.p;
  bool found_zero_pivot = false, ret = true;
  if (size <= 1) {
    if (size == 1 && tid == 0) W.transp[0] = 0;
    block_sync<THREADS>();
    return true;
  }
  for (int kk = 0; kk < size; ++kk) {
    // pivot: largest |diagonal| of the trailing block, first index wins ties (LDLT.h:324-326)
    double key = -1.0;
    int idx = size;
    for (int i = kk + tid; i < size; i += THREADS) {
      const double v = fabs(KM(i, i));
      if (v > key) { key = v; idx = i; }
    }
    const int big = block_argmax<THREADS>(key, idx, W.scratch);
    if (tid == 0) W.transp[kk] = big;
    if (big != kk && big < size) {  // symmetric swap restricted to the lower triangle (LDLT.h:329-345)
      for (int j = tid; j < kk; j += THREADS) { const double t = KM(kk, j); KM(kk, j) = KM(big, j); KM(big, j) = t; }
      for (int i = big + 1 + tid; i < size; i += THREADS) { const double t = KM(i, kk); KM(i, kk) = KM(i, big); KM(i, big) = t; }
      for (int i = kk + 1 + tid; i < big; i += THREADS) { const double t = KM(i, kk); KM(i, kk) = KM(big, i); KM(big, i) = t; }
      if (tid == 0) { const double t = KM(kk, kk); KM(kk, kk) = KM(big, big); KM(big, big) = t; }
    }
    block_sync<THREADS>();
    if (kk > 0) {  // left-looking update of column kk (LDLT.h:356-362)
      for (int j = tid; j < kk; j += THREADS) W.temp[j] = KM(j, j) * KM(kk, j);
      block_sync<THREADS>();
      for (int i = kk + tid; i < size; i += THREADS) {
        double acc = 0.0;
        for (int j = 0; j < kk; ++j) acc += KM(i, j) * W.temp[j];
        KM(i, kk) -= acc;
      }
      block_sync<THREADS>();
    }
    const double akk = KM(kk, kk);
    const bool valid = fabs(akk) > 0.0;
    if (kk == 0 && !valid) {  // whole diagonal is zero (LDLT.h:371-382)
      double bad = 0.0;
      for (int j = 0; j < size; ++j) {
        if (tid == 0) W.transp[j] = j;
        for (int i = j + 1 + tid; i < size; i += THREADS) bad += (KM(i, j) != 0.0) ? 1.0 : 0.0;
      }
      bad = block_sum<THREADS>(bad, W.scratch);
      block_sync<THREADS>();
      return bad == 0.0;
    }
    const int rs = size - kk - 1;
    if (rs > 0) {
      if (valid) {
        for (int i = kk + 1 + tid; i < size; i += THREADS) KM(i, kk) /= akk;
      } else {
        double bad = 0.0;
        for (int i = kk + 1 + tid; i < size; i += THREADS) bad += (KM(i, kk) != 0.0) ? 1.0 : 0.0;
        bad = block_sum<THREADS>(bad, W.scratch);
        ret = ret && (bad == 0.0);
      }
    }
    if (found_zero_pivot && valid) ret = false;  // LDLT.h:389-390
    else if (!valid) found_zero_pivot = true;
    block_sync<THREADS>();
  }
  return ret;
}

// Solve Kmat * x = rhs in place with the factors above (Eigen LDLT::_solve_impl_transposed, LDLT.h:574-611).
template <int THREADS>
__device__ void gen_ldlt_solve(const GenWork& W, double* x) {
  const int tid = team_tid<THREADS>(), size = W.p;
  if (tid == 0) {
    for (int i = 0; i < size; ++i) { const int t = W.transp[i]; if (t != i) { const double v = x[i]; x[i] = x[t]; x[t] = v; } }
  }
  block_sync<THREADS>();
  for (int j = 0; j < size; ++j) {  // unit-lower forward substitution
    const double xj = x[j];
    for (int i = j + 1 + tid; i < size; i += THREADS) x[i] -= KM(i, j) * xj;
    block_sync<THREADS>();
  }
  for (int i = tid; i < size; i += THREADS) {  // pseudo-inverse of D, tolerance DBL_MIN (LDLT.h:595-602)
    const double d = KM(i, i);
    x[i] = (fabs(d) > DBL_MIN) ? x[i] / d : 0.0;
  }
  block_sync<THREADS>();
  for (int j = size - 1; j > 0; --j) {  // L' back substitution, column oriented
    const double xj = x[j];
    for (int i = tid; i < j; i += THREADS) x[i] -= KM(j, i) * xj;
    block_sync<THREADS>();
  }
  if (tid == 0) {
    for (int i = size - 1; i >= 0; --i) { const int t = W.transp[i]; if (t != i) { const double v = x[i]; x[i] = x[t]; x[t] = v; } }
  }
  block_sync<THREADS>();
}


// Carve the workspace out of dynamic shared memory. Same arithmetic on host (size) and device.
__host__ __device__ inline size_t gen_smem_doubles(int n, int k, int m, int threads) {
  const int p = n + k, ld = p | 1, T = n + 2 * m + k;
  size_t d = (size_t)p * ld + 4 * (size_t)T + 2 * (size_t)p + 3 * (size_t)m + 2 * (threads / 32) + 2;
  d += ((size_t)m + p + 1) / 2 + 1;  // int arrays cv[m], transp[p]
  return d;
}

__device__ inline void gen_carve(GenWork& W, double* base, int n, int k, int m, int threads) {
  W.n = n; W.k = k; W.m = m; W.p = n + k; W.ld = W.p | 1; W.T = n + 2 * m + k;
  double* q = base;
  W.Kmat = q; q += (size_t)W.p * W.ld;
  W.vars = q; q += W.T;
  W.r = q; q += W.T;
  W.delta = q; q += W.T;
  W.daff = q; q += W.T;
  W.rhs = q; q += W.p;
  W.temp = q; q += W.p;
  W.w = q; q += m;
  W.ca = q; q += m;
  W.cb = q; q += m;
  W.scratch = q; q += 2 * (threads / 32) + 2;
  W.cv = reinterpret_cast<int*>(q);
  W.transp = W.cv + m;
}

// Pointers of problem `prob` in the batch arrays.
struct GenProb {
  const double* G; const double* c; const double* A; const double* b;
};

#define GX(W) ((W).vars)
#define GS(W) ((W).vars + (W).n)
#define GY(W) ((W).vars + (W).n + (W).m)
#define GZ(W) ((W).vars + (W).n + (W).m + (W).k)

// EvaluateKKTConditions, qp.cc:391-420. The symmetric product reads only the lower triangle of G.
template <int THREADS>
__device__ void gen_eval_kkt(const GenWork& W, const GenProb& Q, bool with_ineq) {
  const int tid = team_tid<THREADS>(), n = W.n, k = W.k, m = W.m;
  const double *x = GX(W), *s = GS(W), *y = GY(W), *z = GZ(W);
  double *r_d = W.r, *r_comp = W.r + n, *r_pe = W.r + n + m, *r_pi = W.r + n + m + k;
  for (int i = tid; i < n; i += THREADS) {
    double acc = 0.0;
    for (int j = 0; j < n; ++j) acc += (i >= j ? Q.G[(size_t)j * n + i] : Q.G[(size_t)i * n + j]) * x[j];
    acc += Q.c[i];
    if (k > 0) {
      double t = 0.0;
      for (int q = 0; q < k; ++q) t += Q.A[(size_t)i * k + q] * y[q];
      acc -= t;
    }
    r_d[i] = acc;
  }
  for (int q = tid; q < k; q += THREADS) {
    double acc = 0.0;
    for (int j = 0; j < n; ++j) acc += Q.A[(size_t)j * k + q] * x[j];
    r_pe[q] = acc + Q.b[q];
  }
  block_sync<THREADS>();
  if (with_ineq) {
    for (int i = tid; i < m; i += THREADS) {
      r_pi[i] = W.ca[i] * x[W.cv[i]] + W.cb[i] - s[i];
      r_comp[i] = s[i] * z[i];
    }
    // r_d[v] -= a * z, in list order per variable (several constraints may share a variable)
    for (int v = tid; v < n; v += THREADS) {
      double acc = r_d[v];
      for (int i = 0; i < m; ++i)
        if (W.cv[i] == v) acc -= W.ca[i] * z[i];
      r_d[v] = acc;
    }
    block_sync<THREADS>();
  }
}

// ComputeLDLT, qp.cc:275-316: assemble [[G + sum a^2 z/s, .],[A, 0]] (lower triangle) and factor.
// Returns 0, MO_QP_NONPOSITIVE_SLACK or MO_QP_FAILED_FACTORIZATION (uniform).
template <int THREADS>
__device__ int gen_compute_ldlt(const GenWork& W, const GenProb& Q, bool with_ineq) {
  const int tid = team_tid<THREADS>(), n = W.n, k = W.k, m = W.m, p = W.p;
  const double *s = GS(W), *z = GZ(W);
  if (with_ineq) {
    double bad = 0.0;
    for (int i = tid; i < m; i += THREADS) bad += (s[i] > 0.0) ? 0.0 : 1.0;
    bad = block_sum<THREADS>(bad, W.scratch);
    if (bad != 0.0) return MO_QP_NONPOSITIVE_SLACK;  // F_ASSERT, qp.cc:285
  }
  for (int e = tid; e < p * p; e += THREADS) {
    const int j = e / p, i = e - j * p;
    double v = 0.0;
    if (j < n) {
      if (i < n) v = (i >= j) ? Q.G[(size_t)j * n + i] : 0.0;
      else v = Q.A[(size_t)j * k + (i - n)];
    }
    KM(i, j) = v;
  }
  block_sync<THREADS>();
  if (with_ineq) {
    for (int i = tid; i < m; i += THREADS) W.w[i] = z[i] / s[i];
    block_sync<THREADS>();
    for (int v = tid; v < n; v += THREADS) {
      double h = KM(v, v);
      for (int i = 0; i < m; ++i)
        if (W.cv[i] == v) h += W.ca[i] * W.w[i] * W.ca[i];
      KM(v, v) = h;
    }
    block_sync<THREADS>();
  }
  if (!gen_ldlt_factor<THREADS>(W)) return MO_QP_FAILED_FACTORIZATION;
  for (int i = tid; i < W.T; i += THREADS) { W.delta[i] = 0.0; W.daff[i] = 0.0; }
  block_sync<THREADS>();
  return 0;
}

// SolveForUpdate (qp.cc:318-364) / SolveForUpdateNoInequalities (:366-386): delta <- Newton direction.
template <int THREADS>
__device__ void gen_solve_for_update(const GenWork& W, double mu, bool with_ineq) {
  const int tid = team_tid<THREADS>(), n = W.n, k = W.k, m = W.m;
  const double *s = GS(W);
  const double *r_d = W.r, *r_comp = W.r + n, *r_pe = W.r + n + m, *r_pi = W.r + n + m + k;
  const double *ds_aff = W.daff + n, *dz_aff = W.daff + n + m + k;
  for (int v = tid; v < n; v += THREADS) {
    double acc = r_d[v];
    if (with_ineq) {
      for (int i = 0; i < m; ++i)
        if (W.cv[i] == v) {
          acc += W.ca[i] * W.w[i] * r_pi[i];
          acc += W.ca[i] * (r_comp[i] + (ds_aff[i] * dz_aff[i]) - mu) / s[i];
        }
    }
    W.rhs[v] = -acc;
  }
  for (int q = tid; q < k; q += THREADS) W.rhs[n + q] = -r_pe[q];
  block_sync<THREADS>();
  gen_ldlt_solve<THREADS>(W, W.rhs);
  double *dx = W.delta, *dS = W.delta + n, *dy = W.delta + n + m, *dz = W.delta + n + m + k;
  for (int i = tid; i < n; i += THREADS) dx[i] = W.rhs[i];
  for (int q = tid; q < k; q += THREADS) dy[q] = -W.rhs[n + q];
  if (with_ineq) {
    for (int i = tid; i < m; i += THREADS) {
      const double d = W.ca[i] * W.rhs[W.cv[i]] + r_pi[i];
      dS[i] = d;
      dz[i] = -W.w[i] * d - (1 / s[i]) * (r_comp[i] + (ds_aff[i] * dz_aff[i]) - mu);
    }
  }
  block_sync<THREADS>();
}


// ComputeAlpha, qp.cc:491-507 (exact formula; the min is order independent).
template <int THREADS>
__device__ double gen_compute_alpha(const GenWork& W, const double* val, const double* dval, double tau) {
  double alpha = 1.0;
  for (int i = team_tid<THREADS>(); i < W.m; i += THREADS) {
    const double updated = val[i] + dval[i];
    if (updated <= 0.0 && fabs(dval[i]) > 0) alpha = fmin(alpha, -tau * val[i] / dval[i]);
  }
  return block_min<THREADS>(alpha, W.scratch);
}

// ComputeMu, qp.cc:509-516
template <int THREADS>
__device__ double gen_compute_mu(const GenWork& W) {
  if (W.m == 0) return 0.0;
  const double *s = GS(W), *z = GZ(W);
  double acc = 0.0;
  for (int i = team_tid<THREADS>(); i < W.m; i += THREADS) acc += s[i] * z[i];
  return block_sum<THREADS>(acc, W.scratch) / (double)W.m;
}

// ComputeErrors, qp.cc:423-437 -> {r_dual, r_comp, r_primal_eq, r_primal_ineq}
template <int THREADS>
__device__ void gen_compute_errors(const GenWork& W, double mu, double (&out)[4]) {
  const int tid = team_tid<THREADS>(), n = W.n, k = W.k, m = W.m;
  const double *r_d = W.r, *r_comp = W.r + n, *r_pe = W.r + n + m, *r_pi = W.r + n + m + k;
  double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
  for (int i = tid; i < n; i += THREADS) a0 += r_d[i] * r_d[i];
  for (int i = tid; i < k; i += THREADS) a1 += r_pe[i] * r_pe[i];
  for (int i = tid; i < m; i += THREADS) { a2 += r_pi[i] * r_pi[i]; a3 += r_comp[i] * r_comp[i]; a4 += r_comp[i]; }
  a0 = block_sum<THREADS>(a0, W.scratch);
  out[0] = sqrt(a0); out[1] = 0.0; out[2] = 0.0; out[3] = 0.0;
  if (k > 0) out[2] = sqrt(block_sum<THREADS>(a1, W.scratch));
  if (m > 0) {
    a2 = block_sum<THREADS>(a2, W.scratch);
    a3 = block_sum<THREADS>(a3, W.scratch);
    a4 = block_sum<THREADS>(a4, W.scratch);
    const double corrected = a3 - 2 * (a4 * mu) + (mu * mu) * m;
    out[1] = sqrt(corrected > 0. ? corrected : 0.);
    out[3] = sqrt(a2);
  }
}

__device__ __forceinline__ double max4(const double (&e)[4]) { return fmax(fmax(e[0], e[1]), fmax(e[2], e[3])); }

// Iterate, qp.cc:153-201. ip = {mu, alpha.primal, alpha.dual, probe.primal, probe.dual, mu_affine}.
// Returns 0 or a failure status (state untouched on failure).
template <int THREADS>
__device__ int gen_iterate(const GenWork& W, const GenProb& Q, double mu_input, int strategy, double (&ip)[6]) {
  const int tid = team_tid<THREADS>(), n = W.n, k = W.k, m = W.m;
  gen_eval_kkt<THREADS>(W, Q, true);
  ip[0] = mu_input; ip[1] = 1.0; ip[2] = 1.0; ip[3] = qnan(); ip[4] = qnan(); ip[5] = qnan();
  const int st = gen_compute_ldlt<THREADS>(W, Q, true);
  if (st) return st;
  double *s = GS(W), *z = GZ(W);
  double *dS = W.delta + n, *dz = W.delta + n + m + k;
  if (m == 0) {
    gen_solve_for_update<THREADS>(W, 0.0, true);
  } else if (strategy != MO_BARRIER_PREDICTOR_CORRECTOR) {
    gen_solve_for_update<THREADS>(W, ip[0], true);
  } else {
    gen_solve_for_update<THREADS>(W, 0.0, true);
    ip[3] = gen_compute_alpha<THREADS>(W, s, dS, 1.0);
    ip[4] = gen_compute_alpha<THREADS>(W, z, dz, 1.0);
    for (int i = tid; i < W.T; i += THREADS) W.daff[i] = W.delta[i];
    block_sync<THREADS>();
    // ComputePredictorCorrectorMuAffine, qp.cc:519-537
    double sdz = 0, zds = 0, dsdz = 0;
    for (int i = tid; i < m; i += THREADS) { sdz += s[i] * dz[i]; zds += z[i] * dS[i]; dsdz += dS[i] * dz[i]; }
    sdz = block_sum<THREADS>(sdz, W.scratch);
    zds = block_sum<THREADS>(zds, W.scratch);
    dsdz = block_sum<THREADS>(dsdz, W.scratch);
    double mu_aff = ip[0];
    mu_aff += ip[4] * sdz / (double)m;
    mu_aff += ip[3] * zds / (double)m;
    mu_aff += (ip[4] * ip[3]) * dsdz / (double)m;
    ip[5] = mu_aff > 0.0 ? mu_aff : 0.0;
    const double r = ip[5] / ip[0];
    ip[0] = (r * r * r) * mu_input;
    gen_solve_for_update<THREADS>(W, ip[0], true);
  }
  if (m > 0) {
    ip[1] = gen_compute_alpha<THREADS>(W, s, dS, 0.995);
    ip[2] = gen_compute_alpha<THREADS>(W, z, dz, 0.995);
  }
  double *x = GX(W), *y = GY(W);
  const double *dx = W.delta, *dy = W.delta + n + m;
  for (int i = tid; i < n; i += THREADS) x[i] += dx[i] * ip[1];
  for (int i = tid; i < m; i += THREADS) { s[i] += dS[i] * ip[1]; z[i] += dz[i] * ip[2]; }
  for (int i = tid; i < k; i += THREADS) y[i] += dy[i] * ip[2];
  block_sync<THREADS>();
  return 0;
}


// Bind problem `prob`: workspace sizes for its own constraint count, constraint list into shared memory.
template <int THREADS>
__device__ void gen_bind(const QPArgs& a, long long prob, double* smem, GenWork& W, GenProb& Q) {
  const int n = a.n, k = a.k, mmax = a.m;
  const int m = a.icount ? a.icount[prob] : mmax;
  gen_carve(W, smem, n, k, m, THREADS);
  Q.G = a.G + prob * (long long)n * n;
  Q.c = a.c + prob * (long long)n;
  Q.A = k ? a.A + prob * (long long)k * n : nullptr;
  Q.b = k ? a.b + prob * (long long)k : nullptr;
  const bool box = (a.flags & MO_QP_FLAG_BOX_PAIRS) != 0;
  for (int i = team_tid<THREADS>(); i < m; i += THREADS) {
    W.cv[i] = box ? i / 2 : a.ivar[prob * mmax + i];
    W.ca[i] = box ? ((i & 1) ? -1.0 : 1.0) : a.ia[prob * mmax + i];
    W.cb[i] = a.ib[prob * mmax + i];
  }
  block_sync<THREADS>();
}

// compact [x|s(m)|y|z(m)] in shared memory <-> fixed stride [x|s(mmax)|y|z(mmax)] in global (unused = NaN)
template <int THREADS>
__device__ void gen_store_vars(const GenWork& W, int mmax, const double* v, double* dst) {
  const int n = W.n, k = W.k, m = W.m;
  for (int i = team_tid<THREADS>(); i < n; i += THREADS) dst[i] = v[i];
  for (int i = team_tid<THREADS>(); i < mmax; i += THREADS) {
    dst[n + i] = i < m ? v[n + i] : qnan();
    dst[n + mmax + k + i] = i < m ? v[n + m + k + i] : qnan();
  }
  for (int i = team_tid<THREADS>(); i < k; i += THREADS) dst[n + mmax + i] = v[n + m + i];
}

template <int THREADS>
__device__ void gen_load_vars(const GenWork& W, int mmax, const double* src, double* v) {
  const int n = W.n, k = W.k, m = W.m;
  for (int i = team_tid<THREADS>(); i < n; i += THREADS) v[i] = src[i];
  for (int i = team_tid<THREADS>(); i < m; i += THREADS) { v[n + i] = src[n + i]; v[n + m + k + i] = src[n + mmax + k + i]; }
  for (int i = team_tid<THREADS>(); i < k; i += THREADS) v[n + m + i] = src[n + mmax + i];
  block_sync<THREADS>();
}

// ComputeInitialGuess, qp.cc:439-482. Returns 0 or a failure status.
template <int THREADS>
__device__ int gen_initial_guess(const GenWork& W, const GenProb& Q, int method) {
  const int tid = team_tid<THREADS>(), n = W.n, k = W.k, m = W.m;
  if (method == MO_GUESS_USER_PROVIDED) return 0;
  double *x = GX(W), *s = GS(W), *y = GY(W), *z = GZ(W);
  for (int i = tid; i < n; i += THREADS) x[i] = 0.0;
  for (int i = tid; i < k; i += THREADS) y[i] = 0.0;
  block_sync<THREADS>();
  if (method == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED) {
    const int st = gen_compute_ldlt<THREADS>(W, Q, false);
    if (st) return st;
    gen_eval_kkt<THREADS>(W, Q, false);
    gen_solve_for_update<THREADS>(W, 0.0, false);
    for (int i = tid; i < n; i += THREADS) x[i] = W.delta[i];
    for (int i = tid; i < k; i += THREADS) y[i] = W.delta[n + m + i];
    block_sync<THREADS>();
  }
  if (tid == 0) {  // sequential clamp in list order (qp.cc:464-467, ClampX qp.hpp:43-53)
    for (int i = 0; i < m; ++i) {
      const int v = W.cv[i];
      const double ca = W.ca[i], cb = W.cb[i], xv = x[v];
      if (ca < 0) { const double hi = cb / -ca; x[v] = xv < hi ? xv : hi; }
      else { const double lo = -cb / ca; x[v] = xv > lo ? xv : lo; }
    }
  }
  block_sync<THREADS>();
  for (int i = tid; i < m; i += THREADS) {  // qp.cc:472-481
    const double sv = W.ca[i] * x[W.cv[i]] + W.cb[i];
    s[i] = fmax(1.0e-9, sv);
    z[i] = 1.0 / s[i];
  }
  block_sync<THREADS>();
  return 0;
}

// Parameters of one Solve() (QPInteriorPointSolver::Params, qp.hpp:134-164).
struct GenSolveParams {
  double initial_mu, sigma, kkt_tol, comp_tol;
  int max_iterations, strategy, decrease_mu_only_on_small_error, guess, init_mu_with_comp;
};

// Solve (qp.cc:100-151) on a bound problem whose W.vars is zeroed (or holds the user's guess). Returns the
// termination status (uniform); `iters` = iterations.size(), `kkt_after` = kkt_final of the last iteration.
// `log` = this problem's iter_log rows or nullptr.
template <int THREADS>
__device__ int gen_solve(const GenWork& W, const GenProb& Q, const GenSolveParams& a, double* log, int& iters,
                         double (&kkt_after)[4]) {
  const int tid = team_tid<THREADS>();
  int status = gen_initial_guess<THREADS>(W, Q, a.guess);
  iters = 0;
  double kkt_prev[4] = {0, 0, 0, 0}, ip[6];
  kkt_after[0] = kkt_after[1] = kkt_after[2] = kkt_after[3] = 0.0;
  if (status) return status;
  status = MO_QP_MAX_ITERATIONS;
  gen_eval_kkt<THREADS>(W, Q, true);
  double mu = a.init_mu_with_comp ? gen_compute_mu<THREADS>(W) : a.initial_mu;
  for (int iter = 0; iter < a.max_iterations; ++iter) {
    gen_compute_errors<THREADS>(W, mu, kkt_prev);
    const int st = gen_iterate<THREADS>(W, Q, mu, a.strategy, ip);
    if (st) { status = st; break; }
    gen_eval_kkt<THREADS>(W, Q, true);
    gen_compute_errors<THREADS>(W, mu, kkt_after);
    if (log && tid == 0) {
      double* row = log + iter * MO_QP_ITER_LOG_STRIDE;
      for (int q = 0; q < 4; ++q) { row[q] = kkt_prev[q]; row[4 + q] = kkt_after[q]; }
      for (int q = 0; q < 6; ++q) row[8 + q] = ip[q];
    }
    ++iters;
    const double mu_c = gen_compute_mu<THREADS>(W);
    if (max4(kkt_after) < a.kkt_tol && mu_c < a.comp_tol) { status = MO_QP_SATISFIED_KKT_TOL; break; }
    if (max4(kkt_after) <= mu || !a.decrease_mu_only_on_small_error) {
      if (a.strategy == MO_BARRIER_FIXED_DECREASE) mu *= a.sigma;
      else mu = a.sigma * mu_c;
    }
  }
  return status;
}

// Solve (qp.cc:100-151): one block per problem, grid-stride over the batch.
template <int THREADS>
__global__ void __launch_bounds__(THREADS) qp_ip_generic_kernel(const QPArgs a) {
  extern __shared__ double smem[];
  const int tid = team_tid<THREADS>();
  const long long vstride = a.n + 2LL * a.m + a.k;
  const GenSolveParams sp{a.initial_mu, a.sigma, a.kkt_tol, a.comp_tol, a.max_iterations, a.strategy,
                          a.decrease_mu_only_on_small_error, a.guess, a.init_mu_with_comp};
  for (long long prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
    GenWork W;
    GenProb Q;
    block_sync<THREADS>();
    gen_bind<THREADS>(a, prob, smem, W, Q);
    if (a.guess == MO_GUESS_USER_PROVIDED) gen_load_vars<THREADS>(W, a.m, a.vars + prob * vstride, W.vars);
    else { for (int i = tid; i < W.T; i += THREADS) W.vars[i] = 0.0; block_sync<THREADS>(); }
    int iters = 0;
    double kkt_after[4];
    const int status = gen_solve<THREADS>(W, Q, sp, a.iter_log ? a.iter_log + prob * a.max_iterations * MO_QP_ITER_LOG_STRIDE : nullptr,
                                          iters, kkt_after);
    block_sync<THREADS>();
    gen_store_vars<THREADS>(W, a.m, W.vars, a.vars + prob * vstride);
    if (tid == 0) {
      a.status[prob] = status;
      a.iters[prob] = iters;
      if (a.kkt_final) for (int q = 0; q < 4; ++q) a.kkt_final[prob * 4 + q] = kkt_after[q];
      if (a.lagrange) {  // ComputeLagrangeMultiplierSummary, qp.cc:539-546
        double mn = qnan(), li = qnan();
        if (W.k > 0) {
          const double* y = GY(W);
          mn = y[0]; li = fabs(y[0]);
          for (int i = 1; i < W.k; ++i) { mn = fmin(mn, y[i]); li = fmax(li, fabs(y[i])); }
        }
        a.lagrange[prob * 2] = mn;
        a.lagrange[prob * 2 + 1] = li;
      }
    }
  }
}

// One Iterate() per problem on caller state (mo_qp_ip_newton_step_batch).
template <int THREADS>
__global__ void __launch_bounds__(THREADS) qp_ip_generic_step_kernel(const QPArgs a, const StepArgs sa) {
  extern __shared__ double smem[];
  const int tid = team_tid<THREADS>();
  const long long vstride = a.n + 2LL * a.m + a.k;
  for (long long prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
    GenWork W;
    GenProb Q;
    block_sync<THREADS>();
    gen_bind<THREADS>(a, prob, smem, W, Q);
    gen_load_vars<THREADS>(W, a.m, a.vars + prob * vstride, W.vars);
    double ip[6];
    const int st = gen_iterate<THREADS>(W, Q, sa.mu, a.strategy, ip);
    if (tid == 0) a.status[prob] = st;
    if (!st) {
      gen_store_vars<THREADS>(W, a.m, W.vars, a.vars + prob * vstride);
      if (sa.delta_out) gen_store_vars<THREADS>(W, a.m, W.delta, sa.delta_out + prob * vstride);
      if (sa.ip_out && tid == 0) for (int q = 0; q < 6; ++q) sa.ip_out[prob * 6 + q] = ip[q];
      if (sa.kkt_out) {
        double e[4];
        gen_eval_kkt<THREADS>(W, Q, true);
        gen_compute_errors<THREADS>(W, sa.mu, e);
        if (tid == 0) for (int q = 0; q < 4; ++q) sa.kkt_out[prob * 4 + q] = e[q];
      }
    }
  }
}

#undef KM

}  // namespace mo
