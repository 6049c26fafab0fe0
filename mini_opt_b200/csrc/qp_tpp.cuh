// Thread-per-problem interior point QP solve for small box-constrained QPs (n + k <= ~14).
//
// Replaces, for every problem of a batch: QPInteriorPointSolver::Setup + Solve
// (source/qp.cc:20-73, 100-151), i.e. ComputeInitialGuess (:439-482), the Iterate loop (:153-201),
// EvaluateKKTConditions (:391-420), ComputeErrors (:423-437), ComputeAlpha (:485-507), ComputeMu (:509-516).
//
// Mapping. One THREAD owns one QP. Its reduced KKT matrix [[G+Sigma, A'],[A, 0]] lives entirely in
// registers as a packed lower triangle and is factored by a fully unrolled LDL' without pivoting
// (the H block is SPD, the Schur block negative definite: quasi-definite, so any order is stable);
// the two triangular solves recover [dx; -dy]. The problem data (G lower triangle, A_eq, c, b_eq, box
// offsets) sits in shared memory, element-major / thread-minor, so every LDS is conflict free.
// Iteration counts differ per problem (7..30), so a warp is a pool of 32 lanes over a contiguous chunk
// of the batch: a lane that terminates stores its result and is REFILLED with the next problem of the
// chunk; the warp loads that problem cooperatively (coalesced 256 B requests) and transposes it into
// the lane's shared-memory column. The equality-constrained initial guess is the same
// assemble/factor/solve code with Sigma = 0, so refilled lanes (phase GUESS) and iterating lanes
// (phase ITER) share the dominant part of the loop body.
#pragma once

#include "qp_common.cuh"

namespace mo {

template <int N, int K>
struct TppLayout {
  static constexpr int P = N + K;
  static constexpr int M = 2 * N;
  static constexpr int NG = N * (N + 1) / 2;
  static constexpr int NK = P * (P + 1) / 2;
  // shared memory element offsets (per thread column)
  static constexpr int OG = 0;
  static constexpr int OA = OG + NG;
  static constexpr int OC = OA + N * K;
  static constexpr int OB = OC + N;
  static constexpr int OIB = OB + K;
  static constexpr int E = OIB + M;
  static constexpr int STRIDE = N + 2 * M + K;  // [x|s|y|z] of the output
};

__host__ __device__ constexpr int lidx(int i, int j) { return i * (i + 1) / 2 + j; }  // i >= j

enum { PHASE_DEAD = 0, PHASE_GUESS = 1, PHASE_ITER = 2 };

// In-register LDL' of the packed lower triangle Km (P x P), natural order, no pivoting, followed by
// the solve of Km * sol = rhs (in place). Returns false when a pivot is zero or not finite
// (-> MO_QP_FAILED_FACTORIZATION, the batched stand-in for qp.cc:303-307).
template <int P>
__device__ __forceinline__ bool ldlt_solve_inreg(double (&Km)[P * (P + 1) / 2], double (&rhs)[P]) {
  double dinv[P];
  bool ok = true;
#pragma unroll
  for (int j = 0; j < P; ++j) {
    const double d = Km[lidx(j, j)];
    ok = ok && (fabs(d) > 0.0) && (fabs(d) < __longlong_as_double(0x7ff0000000000000LL));
    const double inv = fast_rcp(d);
    dinv[j] = inv;
    double t[P];
#pragma unroll
    for (int i = j + 1; i < P; ++i) t[i] = Km[lidx(i, j)];
#pragma unroll
    for (int i = j + 1; i < P; ++i) {
      const double l = t[i] * inv;
#pragma unroll
      for (int c = j + 1; c <= i; ++c) Km[lidx(i, c)] = fma(-l, t[c], Km[lidx(i, c)]);
      Km[lidx(i, j)] = l;
    }
  }
  // L z = rhs
#pragma unroll
  for (int j = 0; j < P; ++j) {
#pragma unroll
    for (int i = j + 1; i < P; ++i) rhs[i] = fma(-Km[lidx(i, j)], rhs[j], rhs[i]);
  }
#pragma unroll
  for (int j = 0; j < P; ++j) rhs[j] *= dinv[j];
  // L' sol = z
#pragma unroll
  for (int j = P - 1; j >= 0; --j) {
#pragma unroll
    for (int i = j + 1; i < P; ++i) rhs[j] = fma(-Km[lidx(i, j)], rhs[i], rhs[j]);
  }
  return ok;
}

// ComputeAlpha (qp.cc:491-507) for one (val, dval) pair, division free: keeps the pair (bv, bd) whose
// ratio -bv/bd is the smallest so far; the caller finishes with alpha = -tau * bv / bd. Starting from
// (1, -tau) reproduces alpha = 1 when no constraint limits the step.
__device__ __forceinline__ void alpha_accumulate(double val, double dval, double& bv, double& bd) {
  const double updated = val + dval;
  // candidate val/(-dval) < bv/(-bd)  <=>  val*(-bd) < bv*(-dval)   (all four factors positive)
  const double diff = fma(val, -bd, bv * dval);  // val*(-bd) - bv*(-dval)
  const bool take = (updated <= 0.0) && (fabs(dval) > 0.0) && (diff < 0.0);
  bv = take ? val : bv;
  bd = take ? dval : bd;
}

// EvaluateKKTConditions (qp.cc:391-420) for a box-pair problem held in the thread's shared-memory
// column `sm` (stride LD between elements), plus the mu-independent sums ComputeErrors needs
// (qp.cc:423-437): sums = {|r_d|^2, |r_pe|^2, |r_pi|^2, sum (s z)^2, sum s z}.
template <int N, int K, bool WITH_INEQ>
__device__ __forceinline__ void eval_kkt(const double* __restrict__ sm, int LD, const double (&x)[N],
                                         const double (&y)[K > 0 ? K : 1], const double (&s)[2 * N],
                                         const double (&z)[2 * N], double (&rd)[N], double (&rpe)[K > 0 ? K : 1],
                                         double (&sums)[5]) {
  using L = TppLayout<N, K>;
  // r_d = G_sym x + c
#pragma unroll
  for (int i = 0; i < N; ++i) rd[i] = sm[(L::OC + i) * LD];
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      const double g = sm[(L::OG + lidx(i, j)) * LD];
      rd[i] = fma(g, x[j], rd[i]);
      if (j != i) rd[j] = fma(g, x[i], rd[j]);
    }
  }
  // r_d -= A' y ; r_pe = A x + b     (A is K x N column-major: (r, j) at j*K + r)
#pragma unroll
  for (int r = 0; r < K; ++r) rpe[r] = sm[(L::OB + r) * LD];
#pragma unroll
  for (int j = 0; j < N; ++j) {
#pragma unroll
    for (int r = 0; r < K; ++r) {
      const double a = sm[(L::OA + j * K + r) * LD];
      rd[j] = fma(-a, y[r], rd[j]);
      rpe[r] = fma(a, x[j], rpe[r]);
    }
  }
  double npi2 = 0.0, sq = 0.0, sum = 0.0;
  if (WITH_INEQ) {
#pragma unroll
    for (int v = 0; v < N; ++v) {
      rd[v] -= z[2 * v];      // a = +1
      rd[v] += z[2 * v + 1];  // a = -1
      const double p0 = (x[v] + sm[(L::OIB + 2 * v) * LD]) - s[2 * v];
      const double p1 = (sm[(L::OIB + 2 * v + 1) * LD] - x[v]) - s[2 * v + 1];
      npi2 = fma(p0, p0, npi2);
      npi2 = fma(p1, p1, npi2);
      const double c0 = s[2 * v] * z[2 * v], c1 = s[2 * v + 1] * z[2 * v + 1];
      sq = fma(c0, c0, sq);
      sq = fma(c1, c1, sq);
      sum += c0;
      sum += c1;
    }
  }
  double nd2 = 0.0, npe2 = 0.0;
#pragma unroll
  for (int i = 0; i < N; ++i) nd2 = fma(rd[i], rd[i], nd2);
#pragma unroll
  for (int r = 0; r < K; ++r) npe2 = fma(rpe[r], rpe[r], npe2);
  sums[0] = nd2; sums[1] = npe2; sums[2] = npi2; sums[3] = sq; sums[4] = sum;
}

// r_comp^2 of ComputeErrors (qp.cc:430-434): max(0, sum (sz)^2 - 2 mu sum sz + mu^2 M)
__device__ __forceinline__ double comp_err2(double sq, double sum, double mu, int M) {
  const double corrected = sq - 2 * (sum * mu) + (mu * mu) * M;
  return corrected > 0.0 ? corrected : 0.0;
}

// Cooperative load of problem q into shared-memory column `col` (element stride LD) by all 32 lanes:
// coalesced 8-byte requests over the problem's contiguous G, A_eq, c, b_eq, ineq_b slices.
template <int N, int K>
__device__ __forceinline__ void load_problem_coop(const QPArgs& a, long long q, double* __restrict__ col, int LD,
                                                  int lane) {
  using L = TppLayout<N, K>;
  const double* __restrict__ Gp = a.G + q * (N * N);
#pragma unroll
  for (int e0 = 0; e0 < N * N; e0 += 32) {
    const int e = e0 + lane;
    if (e < N * N) {
      const int cc = e / N, rr = e - cc * N;  // column-major (rr, cc)
      if (rr >= cc) col[(L::OG + lidx(rr, cc)) * LD] = __ldg(Gp + e);
    }
  }
  constexpr int REST = N * K + N + K + L::M;
  const double* __restrict__ Ap = a.A + q * (N * K);
  const double* __restrict__ cp = a.c + q * N;
  const double* __restrict__ bp = a.b + q * K;
  const double* __restrict__ ibp = a.ib + q * L::M;
#pragma unroll
  for (int e0 = 0; e0 < REST; e0 += 32) {
    const int e = e0 + lane;
    if (e < REST) {
      double v;
      if (K > 0 && e < N * K) v = __ldg(Ap + e);
      else if (e < N * K + N) v = __ldg(cp + (e - N * K));
      else if (K > 0 && e < N * K + N + K) v = __ldg(bp + (e - N * K - N));
      else v = __ldg(ibp + (e - N * K - N - K));
      col[(L::OA + e) * LD] = v;  // OA, OC, OB, OIB are consecutive
    }
  }
}

template <int N, int K, int WPB, int MINB>
__global__ void __launch_bounds__(32 * WPB, MINB) qp_ip_tpp_kernel(const QPArgs a) {
  using L = TppLayout<N, K>;
  constexpr int P = L::P, M = L::M, LD = 32 * WPB + 1;
  constexpr int KD = K > 0 ? K : 1;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* const smcol = smem + threadIdx.x;
  double* const smwarp = smem + warp * 32;

  // contiguous chunk of the batch owned by this warp
  const long long nwarps = (long long)gridDim.x * WPB, gw = (long long)blockIdx.x * WPB + warp;
  const long long chunk = (a.batch + nwarps - 1) / nwarps;
  long long next = gw * chunk < a.batch ? gw * chunk : a.batch;
  const long long wend = next + chunk < a.batch ? next + chunk : a.batch;

  const double tol2 = a.kkt_tol * a.kkt_tol;
  const double tau = 0.995;  // qp.cc:192

  double x[N], y[KD], s[M], z[M], rd[N], rpe[KD], sums[5];
  double mu = a.initial_mu;
  long long prob = -1;
  int iters = 0, phase = PHASE_DEAD;
  bool want = true;
#pragma unroll
  for (int i = 0; i < N; ++i) { x[i] = 0.0; rd[i] = 0.0; }
#pragma unroll
  for (int i = 0; i < KD; ++i) { y[i] = 0.0; rpe[i] = 0.0; }
#pragma unroll
  for (int i = 0; i < M; ++i) { s[i] = 1.0; z[i] = 1.0; }

  for (;;) {
    // ---------------- refill lanes that finished their problem ----------------
    const unsigned need = __ballot_sync(FULL, want);
    if (need) {
      const int rank = __popc(need & ((1u << lane) - 1u));
      if (want) prob = (next + rank < wend) ? next + rank : -1;
      next += __popc(need);
      unsigned todo = need;
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const long long q = __shfl_sync(FULL, prob, src);
        if (q >= 0) load_problem_coop<N, K>(a, q, smwarp + src, LD, lane);
      }
      __syncwarp();
      if (want) {
        want = false;
        phase = prob >= 0 ? PHASE_GUESS : PHASE_DEAD;
        iters = 0;
        if (prob >= 0) {
          // x = y = 0 (qp.cc:445-446)  =>  r_d = c, r_pe = b_eq for the equality-constrained guess
#pragma unroll
          for (int i = 0; i < N; ++i) { x[i] = 0.0; rd[i] = smcol[(L::OC + i) * LD]; }
#pragma unroll
          for (int r = 0; r < K; ++r) { y[r] = 0.0; rpe[r] = smcol[(L::OB + r) * LD]; }
          if (a.guess == MO_GUESS_USER_PROVIDED) {
            const double* __restrict__ vp = a.vars + prob * L::STRIDE;
#pragma unroll
            for (int i = 0; i < N; ++i) x[i] = vp[i];
#pragma unroll
            for (int i = 0; i < M; ++i) { s[i] = vp[N + i]; z[i] = vp[N + M + K + i]; }
#pragma unroll
            for (int r = 0; r < K; ++r) y[r] = vp[N + M + r];
          }
        }
      }
    }
    if (__all_sync(FULL, phase == PHASE_DEAD)) break;
    const bool it = phase == PHASE_ITER;
    int status = -1;  // -1: keep going

    // ---------------- assemble the reduced KKT system (ComputeLDLT qp.cc:275-297, SolveForUpdate :337-342) ----
    double Km[L::NK], rhs[P], is[M];
#pragma unroll
    for (int i = 0; i < L::NG; ++i) Km[i] = smcol[(L::OG + i) * LD];
#pragma unroll
    for (int r = 0; r < K; ++r) {
#pragma unroll
      for (int j = 0; j < N; ++j) Km[lidx(N + r, j)] = smcol[(L::OA + j * K + r) * LD];
#pragma unroll
      for (int q = 0; q <= r; ++q) Km[lidx(N + r, N + q)] = 0.0;
    }
    bool spos = true;
#pragma unroll
    for (int v = 0; v < N; ++v) {
      const double s0 = s[2 * v], s1 = s[2 * v + 1], z0 = z[2 * v], z1 = z[2 * v + 1];
      spos = spos && (s0 > 0.0) && (s1 > 0.0);  // F_ASSERT(all s > 0), qp.cc:285
      const double i0 = fast_rcp(s0), i1 = fast_rcp(s1);
      is[2 * v] = i0;
      is[2 * v + 1] = i1;
      const double w0 = it ? z0 * i0 : 0.0, w1 = it ? z1 * i1 : 0.0;
      double h = Km[lidx(v, v)];
      h += w0;
      h += w1;
      Km[lidx(v, v)] = h;
      const double p0 = (x[v] + smcol[(L::OIB + 2 * v) * LD]) - s0;       // r_pi, a = +1
      const double p1 = (smcol[(L::OIB + 2 * v + 1) * LD] - x[v]) - s1;   // r_pi, a = -1
      double ra = rd[v];
      ra = fma(w0, p0, ra);
      ra = fma(__dmul_rn(s0, z0) - mu, i0, ra);
      ra = fma(-w1, p1, ra);
      ra = fma(-(__dmul_rn(s1, z1) - mu), i1, ra);
      rhs[v] = it ? -ra : -rd[v];
    }
#pragma unroll
    for (int r = 0; r < K; ++r) rhs[N + r] = -rpe[r];
    if (it && !spos) status = MO_QP_NONPOSITIVE_SLACK;

    // ---------------- factor + solve: rhs <- [dx; -dy] ----------------
    const bool skip_solve = (!it) && (a.guess != MO_GUESS_SOLVE_EQUALITY_CONSTRAINED);
    const bool ok = ldlt_solve_inreg<P>(Km, rhs);
    if (status < 0 && !ok && !skip_solve && phase != PHASE_DEAD) status = MO_QP_FAILED_FACTORIZATION;

    double alpha_p = 1.0, alpha_d = 1.0;
    if (it) {
      // ---------------- recover ds, dz (qp.cc:359-363), step lengths (:485-507), update (:196-199) -------
      double ds[M], dz[M];
      double bvp = 1.0, bdp = -tau, bvd = 1.0, bdd = -tau;
#pragma unroll
      for (int v = 0; v < N; ++v) {
        const double s0 = s[2 * v], s1 = s[2 * v + 1], z0 = z[2 * v], z1 = z[2 * v + 1];
        const double i0 = is[2 * v], i1 = is[2 * v + 1];
        const double w0 = z0 * i0, w1 = z1 * i1;
        const double p0 = (x[v] + smcol[(L::OIB + 2 * v) * LD]) - s0;
        const double p1 = (smcol[(L::OIB + 2 * v + 1) * LD] - x[v]) - s1;
        const double d0 = rhs[v] + p0, d1 = p1 - rhs[v];
        const double e0 = -fma(w0, d0, i0 * (__dmul_rn(s0, z0) - mu));
        const double e1 = -fma(w1, d1, i1 * (__dmul_rn(s1, z1) - mu));
        ds[2 * v] = d0; ds[2 * v + 1] = d1;
        dz[2 * v] = e0; dz[2 * v + 1] = e1;
        alpha_accumulate(s0, d0, bvp, bdp);
        alpha_accumulate(s1, d1, bvp, bdp);
        alpha_accumulate(z0, e0, bvd, bdd);
        alpha_accumulate(z1, e1, bvd, bdd);
      }
      alpha_p = (-tau * bvp) / bdp;
      alpha_d = (-tau * bvd) / bdd;
      if (status < 0) {
#pragma unroll
        for (int i = 0; i < N; ++i) x[i] = fma(rhs[i], alpha_p, x[i]);
#pragma unroll
        for (int r = 0; r < K; ++r) y[r] = fma(-rhs[N + r], alpha_d, y[r]);
#pragma unroll
        for (int i = 0; i < M; ++i) { s[i] = fma(ds[i], alpha_p, s[i]); z[i] = fma(dz[i], alpha_d, z[i]); }
      }
    } else if (status < 0) {
      // ---------------- ComputeInitialGuess tail (qp.cc:455-481) ----------------
      if (a.guess == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED) {
#pragma unroll
        for (int i = 0; i < N; ++i) x[i] = rhs[i];
#pragma unroll
        for (int r = 0; r < K; ++r) y[r] = -rhs[N + r];
      }
      if (a.guess != MO_GUESS_USER_PROVIDED) {
#pragma unroll
        for (int v = 0; v < N; ++v) {
          const double b0 = smcol[(L::OIB + 2 * v) * LD], b1 = smcol[(L::OIB + 2 * v + 1) * LD];
          double xv = x[v];
          xv = xv > -b0 ? xv : -b0;  // ClampX, a = +1: max(x, -b/a)   (qp.hpp:43-53)
          xv = xv < b1 ? xv : b1;    // ClampX, a = -1: min(x, b/-a)
          x[v] = xv;
          const double s0 = fmax(1.0e-9, xv + b0), s1 = fmax(1.0e-9, b1 - xv);
          s[2 * v] = s0; s[2 * v + 1] = s1;
          z[2 * v] = 1.0 / s0; z[2 * v + 1] = 1.0 / s1;
        }
      }
    }

    // ---------------- residual at the new state: kkt_final of this step == kkt_initial of the next ------
    if (status < 0) eval_kkt<N, K, true>(smcol, LD, x, y, s, z, rd, rpe, sums);

    if (status < 0) {
      const double mu_c = sums[4] / (double)M;  // ComputeMu, qp.cc:509-516
      if (it) {
        const double comp2 = comp_err2(sums[3], sums[4], mu, M);
        const double max2 = fmax(fmax(sums[0], comp2), fmax(sums[1], sums[2]));
        if (a.iter_log) {
          double* __restrict__ row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
          row[4] = sqrt(sums[0]); row[5] = sqrt(comp2); row[6] = sqrt(sums[1]); row[7] = sqrt(sums[2]);
          row[8] = mu; row[9] = alpha_p; row[10] = alpha_d;
          row[11] = qnan(); row[12] = qnan(); row[13] = qnan();
        }
        ++iters;
        if (max2 < tol2 && mu_c < a.comp_tol) {  // qp.cc:132-137
          status = MO_QP_SATISFIED_KKT_TOL;
        } else if (iters >= a.max_iterations) {
          status = MO_QP_MAX_ITERATIONS;
        } else if (max2 <= mu * mu || !a.decrease_mu_only_on_small_error) {  // qp.cc:140-146
          mu = (a.strategy == MO_BARRIER_FIXED_DECREASE) ? mu * a.sigma : a.sigma * mu_c;
        }
        if (status >= 0 && a.kkt_final) {
          double* __restrict__ kf = a.kkt_final + prob * 4;
          kf[0] = sqrt(sums[0]); kf[1] = sqrt(comp2); kf[2] = sqrt(sums[1]); kf[3] = sqrt(sums[2]);
        }
      } else if (phase == PHASE_GUESS) {
        mu = a.init_mu_with_comp ? mu_c : a.initial_mu;  // qp.cc:115
        phase = PHASE_ITER;
      }
      if (status < 0 && a.iter_log && phase == PHASE_ITER) {
        // kkt_initial of the coming iteration (ComputeErrors with the updated mu, qp.cc:118)
        double* __restrict__ row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
        row[0] = sqrt(sums[0]); row[1] = sqrt(comp_err2(sums[3], sums[4], mu, M));
        row[2] = sqrt(sums[1]); row[3] = sqrt(sums[2]);
      }
    } else if (a.kkt_final && phase != PHASE_DEAD) {
      // failed before completing a step: last completed iteration's errors are not kept; report zeros
      double* __restrict__ kf = a.kkt_final + prob * 4;
      kf[0] = 0.0; kf[1] = 0.0; kf[2] = 0.0; kf[3] = 0.0;
    }

    // ---------------- finished lanes store their result and ask for a new problem ----------------
    if (status >= 0 && phase != PHASE_DEAD) {
      double* __restrict__ vp = a.vars + prob * L::STRIDE;
      const bool zero = (phase == PHASE_GUESS);  // failure inside the initial guess: nothing meaningful yet
#pragma unroll
      for (int i = 0; i < N; ++i) vp[i] = zero ? 0.0 : x[i];
#pragma unroll
      for (int i = 0; i < M; ++i) { vp[N + i] = zero ? 0.0 : s[i]; vp[N + M + K + i] = zero ? 0.0 : z[i]; }
#pragma unroll
      for (int r = 0; r < K; ++r) vp[N + M + r] = zero ? 0.0 : y[r];
      a.status[prob] = status;
      a.iters[prob] = iters;
      if (a.lagrange) {  // ComputeLagrangeMultiplierSummary, qp.cc:539-546
        double mn = qnan(), li = qnan();
        if (K > 0) {
          mn = y[0]; li = fabs(y[0]);
#pragma unroll
          for (int r = 1; r < K; ++r) { mn = y[r] < mn ? y[r] : mn; li = fabs(y[r]) > li ? fabs(y[r]) : li; }
        }
        a.lagrange[prob * 2] = mn;
        a.lagrange[prob * 2 + 1] = li;
      }
      want = true;
      phase = PHASE_DEAD;
    }
  }
}

}  // namespace mo
