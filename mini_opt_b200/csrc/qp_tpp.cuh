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
  static constexpr int E = OIB + M;           // problem data
  static constexpr int OS = E;                // slacks s (state, written by the owning thread)
  static constexpr int OZ = OS + M;           // multipliers z
  static constexpr int ETOT = OZ + M;
  static constexpr int STRIDE = N + 2 * M + K;  // [x|s|y|z] of the output
};

__host__ __device__ constexpr int lidx(int i, int j) { return i * (i + 1) / 2 + j; }  // i >= j

enum { PHASE_DEAD = 0, PHASE_GUESS = 1, PHASE_ITER = 2 };

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

// In-register LDL' of the packed lower triangle Km (P x P), natural order, no pivoting, followed by
// the solve of Km * sol = rhs (in place). The H block is SPD and the Schur block negative definite
// (quasi-definite matrix), so every order is stable; pivots are stored inverted on the diagonal.
// Returns false when a pivot is zero or not finite (-> MO_QP_FAILED_FACTORIZATION, the batched
// stand-in for the exception of qp.cc:303-307).
template <int P>
__device__ __forceinline__ bool ldlt_solve_inreg(double (&Km)[P * (P + 1) / 2], double (&rhs)[P]) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < P; ++j) {
    const double d = Km[lidx(j, j)];
    // zero, NaN and Inf pivots all fail this test (|d| in (0, DBL_MAX])
    ok = ok && (fabs(d) > 0.0) && (fabs(d) <= 1.7976931348623157e308);
    const double inv = fast_rcp(d);
    Km[lidx(j, j)] = inv;
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
#pragma unroll
  for (int j = 0; j < P; ++j) {  // L w = rhs
#pragma unroll
    for (int i = j + 1; i < P; ++i) rhs[i] = fma(-Km[lidx(i, j)], rhs[j], rhs[i]);
  }
#pragma unroll
  for (int j = 0; j < P; ++j) rhs[j] *= Km[lidx(j, j)];  // D^-1
#pragma unroll
  for (int j = P - 1; j >= 0; --j) {  // L' sol = w
#pragma unroll
    for (int i = j + 1; i < P; ++i) rhs[j] = fma(-Km[lidx(i, j)], rhs[i], rhs[j]);
  }
  return ok;
}

// ComputeAlpha (qp.cc:491-507) for one (val, dval) pair without a division: (bv, bd) is the pair with the
// smallest ratio -bv/bd seen so far, starting from (1, -1). A candidate replaces it when
//   val/(-dval) <= bv/(-bd)   <=>   val*(-bd) + bv*dval <= 0      (val, bv, -bd > 0);
// against the initial pair this is exactly the reference's guard val + dval <= 0 (which also implies
// |dval| > 0 because val > 0). The caller finishes with alpha = any ? -tau*bv/bd : 1.
__device__ __forceinline__ void alpha_accumulate(double val, double dval, double& bv, double& bd, bool& any) {
  const double diff = fma(val, -bd, bv * dval);
  const bool take = diff <= 0.0;
  bv = take ? val : bv;
  bd = take ? dval : bd;
  any = any || take;
}

// r_comp^2 of ComputeErrors (qp.cc:430-434): max(0, sum (sz)^2 - 2 mu sum sz + mu^2 M)
__device__ __forceinline__ double comp_err2(double sq, double sum, double mu, int M) {
  const double corrected = sq - 2 * (sum * mu) + (mu * mu) * M;
  return corrected > 0.0 ? corrected : 0.0;
}

// Store a finished problem: [x | s | y | z], status, iteration count, final KKT norms, multiplier summary.
template <int N, int K>
__device__ __forceinline__ void store_result(const QPArgs& a, long long prob, const double* __restrict__ smcol, int LD,
                                             const double (&x)[N], const double (&y)[K > 0 ? K : 1], int status,
                                             int iters, bool zero_state, const double (&kkt2)[4]) {
  using L = TppLayout<N, K>;
  constexpr int M = L::M;
  double* __restrict__ vp = a.vars + prob * L::STRIDE;
#pragma unroll
  for (int i = 0; i < N; ++i) vp[i] = zero_state ? 0.0 : x[i];
#pragma unroll
  for (int i = 0; i < M; ++i) {
    vp[N + i] = zero_state ? 0.0 : smcol[(L::OS + i) * LD];
    vp[N + M + K + i] = zero_state ? 0.0 : smcol[(L::OZ + i) * LD];
  }
#pragma unroll
  for (int r = 0; r < K; ++r) vp[N + M + r] = zero_state ? 0.0 : y[r];
  a.status[prob] = status;
  a.iters[prob] = iters;
  if (a.kkt_final) {
    double* __restrict__ kf = a.kkt_final + prob * 4;
#pragma unroll
    for (int q = 0; q < 4; ++q) kf[q] = sqrt(kkt2[q]);
  }
  if (a.lagrange) {  // ComputeLagrangeMultiplierSummary, qp.cc:539-546
    double mn = qnan(), li = qnan();
    if (K > 0 && !zero_state) {
      mn = y[0]; li = fabs(y[0]);
#pragma unroll
      for (int r = 1; r < K; ++r) { mn = y[r] < mn ? y[r] : mn; li = fabs(y[r]) > li ? fabs(y[r]) : li; }
    } else if (K > 0) {
      mn = 0.0; li = 0.0;
    }
    a.lagrange[prob * 2] = mn;
    a.lagrange[prob * 2 + 1] = li;
  }
}

// ComputeInitialGuess tail (qp.cc:462-481): clamp x into the box in list order, s = max(1e-9, a x + b), z = 1/s.
template <int N, int K>
__device__ __forceinline__ void clamp_and_init_slacks(double* __restrict__ smcol, int LD, double (&x)[N]) {
  using L = TppLayout<N, K>;
#pragma unroll
  for (int v = 0; v < N; ++v) {
    const double b0 = smcol[(L::OIB + 2 * v) * LD], b1 = smcol[(L::OIB + 2 * v + 1) * LD];
    double xv = x[v];
    xv = xv > -b0 ? xv : -b0;  // ClampX, a = +1: max(x, -b/a)   (qp.hpp:43-53)
    xv = xv < b1 ? xv : b1;    // ClampX, a = -1: min(x, b/-a)
    x[v] = xv;
    const double s0 = fmax(1.0e-9, xv + b0), s1 = fmax(1.0e-9, b1 - xv);
    smcol[(L::OS + 2 * v) * LD] = s0;
    smcol[(L::OS + 2 * v + 1) * LD] = s1;
    smcol[(L::OZ + 2 * v) * LD] = 1.0 / s0;
    smcol[(L::OZ + 2 * v + 1) * LD] = 1.0 / s1;
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
  const bool eq_guess = a.guess == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED;

  double x[N], y[KD];
  double mu = a.initial_mu, alpha_p = 1.0, alpha_d = 1.0;
  long long prob = -1;
  int iters = 0, phase = PHASE_DEAD;
  bool want = true, first = false;
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = 0.0;
#pragma unroll
  for (int i = 0; i < KD; ++i) y[i] = 0.0;

  for (;;) {
    double Km[L::NK], rhs[P];
    bool need_eval = phase == PHASE_ITER;
    // ======== phase 1: evaluate / test termination / refill, until every live lane holds a system ========
    for (;;) {
      if (need_eval) {
        need_eval = false;
        int status = -1;
        // ---- EvaluateKKTConditions (qp.cc:391-420) fused with the assembly of ComputeLDLT (:289-297) ----
        double base[N];  // G_sym x + c - A'y
#pragma unroll
        for (int i = 0; i < N; ++i) base[i] = smcol[(L::OC + i) * LD];
#pragma unroll
        for (int i = 0; i < N; ++i) {
#pragma unroll
          for (int j = 0; j <= i; ++j) {
            const double g = smcol[(L::OG + lidx(i, j)) * LD];
            Km[lidx(i, j)] = g;
            base[i] = fma(g, x[j], base[i]);
            if (j != i) base[j] = fma(g, x[i], base[j]);
          }
        }
        double npe2 = 0.0;
#pragma unroll
        for (int r = 0; r < K; ++r) {
          double pe = smcol[(L::OB + r) * LD];
#pragma unroll
          for (int j = 0; j < N; ++j) {
            const double av = smcol[(L::OA + j * K + r) * LD];  // A is K x N column-major
            Km[lidx(N + r, j)] = av;
            base[j] = fma(-av, y[r], base[j]);
            pe = fma(av, x[j], pe);
          }
#pragma unroll
          for (int q = 0; q <= r; ++q) Km[lidx(N + r, N + q)] = 0.0;
          rhs[N + r] = -pe;
          npe2 = fma(pe, pe, npe2);
        }
        double qv[N], tv[N];
        double nd2 = 0.0, npi2 = 0.0, sq = 0.0, sum = 0.0;
        bool spos = true;
#pragma unroll
        for (int v = 0; v < N; ++v) {
          const double s0 = smcol[(L::OS + 2 * v) * LD], s1 = smcol[(L::OS + 2 * v + 1) * LD];
          const double z0 = smcol[(L::OZ + 2 * v) * LD], z1 = smcol[(L::OZ + 2 * v + 1) * LD];
          spos = spos && (s0 > 0.0) && (s1 > 0.0);  // F_ASSERT(all s > 0), qp.cc:285
          const double p0 = (x[v] + smcol[(L::OIB + 2 * v) * LD]) - s0;      // r_pi, a = +1
          const double p1 = (smcol[(L::OIB + 2 * v + 1) * LD] - x[v]) - s1;  // r_pi, a = -1
          const double i0 = fast_rcp(s0), i1 = fast_rcp(s1);
          const double w0 = z0 * i0, w1 = z1 * i1;  // Sigma = z / s
          Km[lidx(v, v)] = (Km[lidx(v, v)] + w0) + w1;
          const double rdv = (base[v] - z0) + z1;  // r_d
          nd2 = fma(rdv, rdv, nd2);
          npi2 = fma(p0, p0, npi2);
          npi2 = fma(p1, p1, npi2);
          const double c0 = s0 * z0, c1 = s1 * z1;  // r_comp
          sq = fma(c0, c0, sq);
          sq = fma(c1, c1, sq);
          sum += c0;
          sum += c1;
          // r_aug = r_d + sum a [w r_pi + (s z - mu)/s]  (qp.cc:337-342)  ==  base + (w0 p0 - w1 p1) - mu (1/s0 - 1/s1)
          qv[v] = fma(w0, p0, -(w1 * p1));
          tv[v] = i0 - i1;
        }
        // ---- Solve loop bookkeeping (qp.cc:110-146): errors of the step just taken, termination, mu ----
        const double mu_c = sum / (double)M;  // ComputeMu, qp.cc:509-516
        double kkt2[4] = {nd2, 0.0, npe2, npi2};
        if (first) {
          first = false;
          mu = a.init_mu_with_comp ? mu_c : a.initial_mu;  // qp.cc:115
        } else {
          kkt2[1] = comp_err2(sq, sum, mu, M);  // kkt_final of the iteration uses the mu it was taken with
          const double max2 = fmax(fmax(kkt2[0], kkt2[1]), fmax(kkt2[2], kkt2[3]));
          if (a.iter_log) {
            double* __restrict__ row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
            row[4] = sqrt(kkt2[0]); row[5] = sqrt(kkt2[1]); row[6] = sqrt(kkt2[2]); row[7] = sqrt(kkt2[3]);
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
        }
        if (status < 0 && !spos) {
          status = MO_QP_NONPOSITIVE_SLACK;
          kkt2[0] = 0.0; kkt2[1] = 0.0; kkt2[2] = 0.0; kkt2[3] = 0.0;
        }
        if (status >= 0) {
          store_result<N, K>(a, prob, smcol, LD, x, y, status, iters, false, kkt2);
          want = true;
          phase = PHASE_DEAD;
        } else {
          if (a.iter_log) {  // kkt_initial of the coming iteration: ComputeErrors with the updated mu (qp.cc:118)
            double* __restrict__ row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
            row[0] = sqrt(nd2); row[1] = sqrt(comp_err2(sq, sum, mu, M)); row[2] = sqrt(npe2); row[3] = sqrt(npi2);
          }
#pragma unroll
          for (int v = 0; v < N; ++v) rhs[v] = -fma(-mu, tv[v], base[v] + qv[v]);
        }
      }

      // ---- refill lanes whose problem finished (or that never had one) ----
      const unsigned need = __ballot_sync(FULL, want);
      if (need == 0u) break;
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
        iters = 0;
        alpha_p = 1.0; alpha_d = 1.0;
        if (prob < 0) {
          phase = PHASE_DEAD;
        } else if (eq_guess) {
          // equality-constrained guess (qp.cc:452-460): the KKT system without Sigma at x = y = 0,
          // where r_d = c and r_pe = b_eq
          phase = PHASE_GUESS;
#pragma unroll
          for (int i = 0; i < L::NG; ++i) Km[i] = smcol[(L::OG + i) * LD];
#pragma unroll
          for (int r = 0; r < K; ++r) {
#pragma unroll
            for (int j = 0; j < N; ++j) Km[lidx(N + r, j)] = smcol[(L::OA + j * K + r) * LD];
#pragma unroll
            for (int q = 0; q <= r; ++q) Km[lidx(N + r, N + q)] = 0.0;
            rhs[N + r] = -smcol[(L::OB + r) * LD];
          }
#pragma unroll
          for (int i = 0; i < N; ++i) rhs[i] = -smcol[(L::OC + i) * LD];
        } else {
          phase = PHASE_ITER;
          first = true;
          need_eval = true;
          if (a.guess == MO_GUESS_USER_PROVIDED) {
            const double* __restrict__ vp = a.vars + prob * L::STRIDE;
#pragma unroll
            for (int i = 0; i < N; ++i) x[i] = vp[i];
#pragma unroll
            for (int i = 0; i < M; ++i) {
              smcol[(L::OS + i) * LD] = vp[N + i];
              smcol[(L::OZ + i) * LD] = vp[N + M + K + i];
            }
#pragma unroll
            for (int r = 0; r < K; ++r) y[r] = vp[N + M + r];
          } else {  // NAIVE: x = y = 0
#pragma unroll
            for (int i = 0; i < N; ++i) x[i] = 0.0;
#pragma unroll
            for (int r = 0; r < K; ++r) y[r] = 0.0;
            clamp_and_init_slacks<N, K>(smcol, LD, x);
          }
        }
      }
      if (!__any_sync(FULL, need_eval)) break;
    }
    if (__all_sync(FULL, phase == PHASE_DEAD)) break;

    // ======== phase 2: factor the reduced KKT system and solve: rhs <- [dx; -dy] ========
    const bool ok = ldlt_solve_inreg<P>(Km, rhs);

    if (phase != PHASE_DEAD && !ok) {
      // FailedFactorization (qp.cc:303-307) as a per-problem status; the state is left as it was
      const double zero4[4] = {0.0, 0.0, 0.0, 0.0};
      store_result<N, K>(a, prob, smcol, LD, x, y, MO_QP_FAILED_FACTORIZATION, iters, phase == PHASE_GUESS, zero4);
      want = true;
      phase = PHASE_DEAD;
    } else if (phase == PHASE_ITER) {
      // ======== phase 3: ds, dz (qp.cc:359-363), step lengths (:485-507), state update (:196-199) ========
      double sv[M], zv[M], ds[M], dz[M];
      double bvp = 1.0, bdp = -1.0, bvd = 1.0, bdd = -1.0;
      bool anyp = false, anyd = false;
#pragma unroll
      for (int v = 0; v < N; ++v) {
        const double s0 = smcol[(L::OS + 2 * v) * LD], s1 = smcol[(L::OS + 2 * v + 1) * LD];
        const double z0 = smcol[(L::OZ + 2 * v) * LD], z1 = smcol[(L::OZ + 2 * v + 1) * LD];
        const double p0 = (x[v] + smcol[(L::OIB + 2 * v) * LD]) - s0;
        const double p1 = (smcol[(L::OIB + 2 * v + 1) * LD] - x[v]) - s1;
        const double i0 = fast_rcp(s0), i1 = fast_rcp(s1);
        const double d0 = rhs[v] + p0, d1 = p1 - rhs[v];  // ds = a dx + r_pi
        // dz = -(z/s) ds - (s z - mu)/s  =  mu/s - (z/s) ds - z
        const double e0 = fma(mu, i0, -fma(z0 * i0, d0, z0));
        const double e1 = fma(mu, i1, -fma(z1 * i1, d1, z1));
        sv[2 * v] = s0; sv[2 * v + 1] = s1; zv[2 * v] = z0; zv[2 * v + 1] = z1;
        ds[2 * v] = d0; ds[2 * v + 1] = d1; dz[2 * v] = e0; dz[2 * v + 1] = e1;
        alpha_accumulate(s0, d0, bvp, bdp, anyp);
        alpha_accumulate(s1, d1, bvp, bdp, anyp);
        alpha_accumulate(z0, e0, bvd, bdd, anyd);
        alpha_accumulate(z1, e1, bvd, bdd, anyd);
      }
      alpha_p = anyp ? (-tau * bvp) * fast_rcp(bdp) : 1.0;
      alpha_d = anyd ? (-tau * bvd) * fast_rcp(bdd) : 1.0;
#pragma unroll
      for (int i = 0; i < N; ++i) x[i] = fma(rhs[i], alpha_p, x[i]);
#pragma unroll
      for (int r = 0; r < K; ++r) y[r] = fma(-rhs[N + r], alpha_d, y[r]);
#pragma unroll
      for (int i = 0; i < M; ++i) {
        smcol[(L::OS + i) * LD] = fma(ds[i], alpha_p, sv[i]);
        smcol[(L::OZ + i) * LD] = fma(dz[i], alpha_d, zv[i]);
      }
    } else if (phase == PHASE_GUESS) {
      // x, y of the equality-constrained problem, then clamp / slack initialisation (qp.cc:455-481)
#pragma unroll
      for (int i = 0; i < N; ++i) x[i] = rhs[i];
#pragma unroll
      for (int r = 0; r < K; ++r) y[r] = -rhs[N + r];
      clamp_and_init_slacks<N, K>(smcol, LD, x);
      phase = PHASE_ITER;
      first = true;
    }
  }
}

}  // namespace mo
