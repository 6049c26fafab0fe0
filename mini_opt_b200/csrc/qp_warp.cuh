// Warp-synchronous interior point solve for SMALL QPs (n + k <= 16, m <= 32, at most two diagonal constraints per
// variable): the inner solver of the chain-IK kernel. Same Solve / Iterate bookkeeping as gen_solve (qp.cc:100-201,
// 275-364, 391-516), but the whole state lives in registers -- lane L < n+k carries KKT row L and the primal-dual
// entry (x_L or y_{L-n}), lane c < m carries constraint c (s, z, a, b, var) -- and the LDL' is unpivoted, right-looking,
// row-per-lane with shuffles. Why: gen_solve<32> follows Eigen's pivoted LDLT element by element; inlined into the IK
// kernel it is 27K instructions and the kernel spent 76 % of its stall samples on instruction fetch
// (profiles/r01_ik6_ncu_summary.txt). This solver is ~1.5K instructions and ~10x fewer issued per Newton step.
// Arithmetic differs from gen_solve at rounding level only (natural pivot order, tree reductions).
#pragma once
#include "qp_generic.cuh"

namespace mo {

constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ double wip_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}
__device__ __forceinline__ double wip_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(kFull, v, o));
  return v;
}

__host__ __device__ inline bool wip_shape_ok(int n, int k, int m) { return n >= 1 && n + k <= 16 && m <= 32; }

// In-register LDL' of the rows in `a` (lane = row, a[j] = K(lane, j), lower triangle significant) + solve of one
// right-hand side (lane = entry). L is parked in shared memory (Lsm, p x ld column-major like KM) for the backward
// pass. Returns false (uniformly) on a zero or non-finite pivot.
template <int PMAX>
__device__ __forceinline__ bool wip_factor_solve(double (&a)[PMAX], int p, double* __restrict__ Lsm, int ld, double& y) {
  const int lane = threadIdx.x & 31;
  double dmine = 1.0, chk = 0.0;
#pragma unroll
  for (int c = 0; c < PMAX; ++c) {
    if (c < p) {
      const double d = __shfl_sync(kFull, a[c], c);
      const double inv = 1.0 / d;
      chk = fma(inv, 0.0, chk);  // NaN for d = 0 / inf / NaN
      const double raw = a[c], l = raw * inv;
      if (lane == c) dmine = d;
#pragma unroll
      for (int c2 = c + 1; c2 < PMAX; ++c2) {
        if (c2 < p) {
          const double v = __shfl_sync(kFull, raw, c2);
          a[c2] = fma(-l, v, a[c2]);
        }
      }
      a[c] = l;
      if (lane > c && lane < p) Lsm[(size_t)c * ld + lane] = l;
      // forward substitution rides along: y_lane -= L(lane, c) y_c
      const double yc = __shfl_sync(kFull, y, c);
      if (lane > c && lane < p) y = fma(-l, yc, y);
    }
  }
  if (chk != chk) return false;
  y = y / dmine;
  __syncwarp();
  for (int c = p - 1; c > 0; --c) {  // L' back substitution, column oriented
    const double yc = __shfl_sync(kFull, y, c);
    if (lane < c) y = fma(-Lsm[(size_t)lane * ld + c], yc, y);
  }
  return true;
}

// Returns the termination status, or -1 when the problem is outside this solver's envelope (caller falls back to
// gen_solve). On return W.vars = [x|s|y|z]. blockDim.x must be 32.
template <int PMAX>
__device__ int wip_solve(const GenWork& W, const GenProb& Q, const GenSolveParams& prm, int& iters, double (&kkt_after)[4]) {
  const int lane = threadIdx.x & 31, n = W.n, k = W.k, m = W.m, p = n + k;
  iters = 0;
  kkt_after[0] = kkt_after[1] = kkt_after[2] = kkt_after[3] = 0.0;
  if (!wip_shape_ok(n, k, m) || prm.strategy == MO_BARRIER_PREDICTOR_CORRECTOR || prm.guess == MO_GUESS_USER_PROVIDED) return -1;
  const bool isx = lane < n, isy = lane >= n && lane < p, isc = lane < m;
  const double ca = isc ? W.ca[lane] : 0.0, cb = isc ? W.cb[lane] : 0.0;
  const int cv = isc ? W.cv[lane] : 0;
  // variable role: my (at most two) constraints, in list order
  int c0 = -1, c1 = -1;
  bool more = false;
  for (int c = 0; c < m; ++c) {
    const int v = __shfl_sync(kFull, cv, c);
    if (isx && v == lane) {
      if (c0 < 0) c0 = c; else if (c1 < 0) c1 = c; else more = true;
    }
  }
  if (__any_sync(kFull, more)) return -1;
  const int s0 = c0 < 0 ? 0 : c0, s1 = c1 < 0 ? 0 : c1;
  // base + sign * (q[c0] + q[c1]) applied in list order
  auto scatter = [&](double base, double q, double sign) {
    const double q0 = __shfl_sync(kFull, q, s0), q1 = __shfl_sync(kFull, q, s1);
    if (c0 >= 0) base = fma(sign, q0, base);
    if (c1 >= 0) base = fma(sign, q1, base);
    return base;
  };
  // my row of [[G, A'],[A, 0]] (both triangles) and my entry of [c; b]
  double g[PMAX];
#pragma unroll
  for (int j = 0; j < PMAX; ++j) {
    double v = 0.0;
    if (isx) {
      if (j < n) v = lane >= j ? Q.G[(size_t)j * n + lane] : Q.G[(size_t)lane * n + j];
      else if (j < p) v = Q.A[(size_t)lane * k + (j - n)];
    } else if (isy && j < n) {
      v = Q.A[(size_t)j * k + (lane - n)];
    }
    g[j] = v;
  }
  const double lin = isx ? Q.c[lane] : (isy ? Q.b[lane - n] : 0.0);
  double* __restrict__ Lsm = W.Kmat;
  const int ld = W.ld;
  double a[PMAX];
  double u = 0.0;  // x_lane or y_{lane-n}
  double s = 1.0, z = 1.0;
  // ComputeInitialGuess tail (qp.cc:462-481): clamp in list order (ClampX, qp.hpp:43-53), s = max(1e-9, a x + b), z = 1/s
  auto clamp_init = [&]() {
    const double a0 = __shfl_sync(kFull, ca, s0), b0 = __shfl_sync(kFull, cb, s0);
    const double a1 = __shfl_sync(kFull, ca, s1), b1 = __shfl_sync(kFull, cb, s1);
    if (c0 >= 0) { if (a0 < 0) { const double hi = b0 / -a0; u = u < hi ? u : hi; } else { const double lo = -b0 / a0; u = u > lo ? u : lo; } }
    if (c1 >= 0) { if (a1 < 0) { const double hi = b1 / -a1; u = u < hi ? u : hi; } else { const double lo = -b1 / a1; u = u > lo ? u : lo; } }
    const double xv = __shfl_sync(kFull, u, cv);
    if (isc) { s = fmax(1.0e-9, fma(ca, xv, cb)); z = 1.0 / s; }
  };
  // SOLVE_EQUALITY_CONSTRAINED (qp.cc:452-459) shares the factor/solve call site of the Newton loop (guess_pending)
  bool guess_pending = prm.guess == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED;
  if (!guess_pending) clamp_init();
  // ---- Solve loop (qp.cc:110-149): each trip = residuals of the current point, bookkeeping, one Newton step ----
  double mu = prm.initial_mu;
  bool first = true;
  int status = MO_QP_MAX_ITERATIONS;
  const double minv = m > 0 ? 1.0 / (double)m : 0.0;
  for (;;) {
    double y, sig = 0.0, rpi = 0.0, comp = 0.0, w = 0.0, is = 0.0, barrier = 0.0;
    if (guess_pending) {
      y = -lin;  // KKT system without Sigma, rhs = -[c; b_eq]
    } else {
    // EvaluateKKTConditions (qp.cc:391-420)
    double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
    for (int j = 0; j < PMAX; j += 2) {
      if (j < p) { const double t = __shfl_sync(kFull, u, j); acc0 = fma(g[j], j < n ? t : -t, acc0); }
      if (j + 1 < p) { const double t = __shfl_sync(kFull, u, j + 1); acc1 = fma(g[j + 1], j + 1 < n ? t : -t, acc1); }
    }
    const double rmain = (acc0 + acc1) + lin;  // r_d before the inequality term (lanes < n), r_pe (lanes n..p)
    const double xv = __shfl_sync(kFull, u, cv);
    rpi = isc ? fma(ca, xv, cb) - s : 0.0; comp = isc ? s * z : 0.0;
    const double rd_all = scatter(rmain, ca * z, -1.0);  // (shuffles: every lane executes, then select)
    const double rd = isx ? rd_all : 0.0;
    // ComputeErrors (qp.cc:423-437), ComputeMu (:509-516)
    const double e_d = wip_sum(isx ? rd * rd : 0.0), e_pe = wip_sum(isy ? rmain * rmain : 0.0);
    const double e_pi = wip_sum(rpi * rpi), e_c2 = wip_sum(comp * comp), e_c = wip_sum(comp);
    if (first) {
      first = false;
      if (prm.init_mu_with_comp) mu = e_c * minv;
    } else {
      const double corrected = e_c2 - 2 * (e_c * mu) + (mu * mu) * m;
      kkt_after[0] = sqrt(e_d); kkt_after[1] = m > 0 ? sqrt(corrected > 0. ? corrected : 0.) : 0.0;
      kkt_after[2] = k > 0 ? sqrt(e_pe) : 0.0; kkt_after[3] = m > 0 ? sqrt(e_pi) : 0.0;
      ++iters;
      const double mu_c = e_c * minv, mx = max4(kkt_after);
      if (mx < prm.kkt_tol && mu_c < prm.comp_tol) { status = MO_QP_SATISFIED_KKT_TOL; break; }
      if (iters >= prm.max_iterations) break;
      if (mx <= mu || !prm.decrease_mu_only_on_small_error)
        mu = prm.strategy == MO_BARRIER_FIXED_DECREASE ? mu * prm.sigma : prm.sigma * mu_c;
    }
    if (prm.max_iterations <= 0) break;
    // ComputeLDLT (qp.cc:275-316)
    if (__any_sync(kFull, isc && !(s > 0.0))) { status = MO_QP_NONPOSITIVE_SLACK; break; }
    w = isc ? z / s : 0.0; is = isc ? 1.0 / s : 0.0;
    barrier = m > 0 ? mu : 0.0;  // Iterate solves with mu = 0 when there are no inequalities
    sig = scatter(0.0, (ca * w) * ca, 1.0);  // 0 on lanes without constraints
    // SolveForUpdate (qp.cc:318-364)
    const double q = fma(ca * w, rpi, ca * (comp - barrier) * is);
    const double raug = scatter(rd, q, 1.0);
    y = isx ? -raug : (isy ? -rmain : 0.0);
    }
#pragma unroll
    for (int j = 0; j < PMAX; ++j) a[j] = g[j] + ((j == lane && isx) ? sig : 0.0);
    if (!wip_factor_solve<PMAX>(a, p, Lsm, ld, y)) { status = MO_QP_FAILED_FACTORIZATION; break; }
    if (guess_pending) {
      guess_pending = false;
      u = isx ? y : (isy ? -y : 0.0);
      clamp_init();
      continue;
    }
    const double dxv = __shfl_sync(kFull, y, cv);
    const double ds = isc ? fma(ca, dxv, rpi) : 0.0;
    const double dz = isc ? -w * ds - is * (comp - barrier) : 0.0;
    // ComputeAlpha (qp.cc:485-507), tau = 0.995 (qp.cc:192)
    double ap = 1.0, ad = 1.0;
    if (isc && s + ds <= 0.0 && fabs(ds) > 0) ap = -0.995 * s / ds;
    if (isc && z + dz <= 0.0 && fabs(dz) > 0) ad = -0.995 * z / dz;
    ap = wip_min(ap); ad = wip_min(ad);
    u = isx ? fma(y, ap, u) : (isy ? fma(-y, ad, u) : 0.0);
    s = fma(ds, ap, s); z = fma(dz, ad, z);
  }
  __syncwarp();
  if (isx) W.vars[lane] = u;
  if (isy) W.vars[n + m + (lane - n)] = u;
  if (isc) { W.vars[n + lane] = s; W.vars[n + m + k + lane] = z; }
  __syncwarp();
  return status;
}

// Smallest register tile that holds the problem: the unrolled column loops carry one uniform branch per column of
// PMAX, so a 6x6 system runs through the PMAX = 8 instance, not through 16 columns of skipped code.
__device__ __forceinline__ int wip_solve_any(const GenWork& W, const GenProb& Q, const GenSolveParams& prm, int& iters,
                                             double (&kkt_after)[4]) {
  const int p = W.n + W.k;
  if (p <= 4) return wip_solve<4>(W, Q, prm, iters, kkt_after);
  if (p <= 8) return wip_solve<8>(W, Q, prm, iters, kkt_after);
  if (p <= 12) return wip_solve<12>(W, Q, prm, iters, kkt_after);
  return wip_solve<16>(W, Q, prm, iters, kkt_after);
}

}  // namespace mo
