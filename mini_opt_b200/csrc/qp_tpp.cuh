// Thread-per-problem interior point QP solve for small box-constrained QPs (n + k <= ~14).
//
// Replaces, for every problem of a batch: QPInteriorPointSolver::Setup + Solve
// (source/qp.cc:20-73, 100-151), i.e. ComputeInitialGuess (:439-482), the Iterate loop (:153-201),
// EvaluateKKTConditions (:391-420), ComputeErrors (:423-437), ComputeAlpha (:485-507), ComputeMu (:509-516).
//
// Mapping. One THREAD owns one QP. Its reduced KKT matrix [[G+Sigma, A'],[A, 0]] lives entirely in
// registers as a packed lower triangle and is factored by a fully unrolled LDL' without pivoting
// (the H block is SPD, the Schur block negative definite: quasi-definite, so any order is stable);
// two triangular solves recover [dx; -dy]. Problem data (G lower triangle, A_eq, c, b_eq, box offsets) and
// the slacks / multipliers s, z sit in shared memory, element-major / thread-minor, so every LDS/STS is
// conflict free; x and y stay in registers.
//
// Two kernels:
//   qp_tpp_guess_kernel    ComputeInitialGuess for every problem, no divergence: writes the start state
//                          [x|s|y|z] into `vars` (the buffer the result will overwrite).
//   qp_tpp_iterate_kernel  the Newton loop from the state in `vars`. Iteration counts differ per problem
//                          (7..30), so a warp is a pool of 32 lanes over a contiguous chunk of the batch: a lane
//                          whose problem terminated parks its x, y in its shared-memory column, the warp
//                          writes the 336-byte result with coalesced stores, and the lane is REFILLED with
//                          the next problem of the chunk (problem + start state loaded cooperatively with
//                          coalesced requests, transposed into the lane's column; the lines a warp needs
//                          next are prefetched into L2 a few problems ahead).
// Each round of the loop is: refill -> residuals + KKT assembly -> termination test / mu update -> LDL' +
// solve -> step lengths + state update; the only per-lane divergence is "this lane's problem just ended".
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

// Cooperative load of `cnt` (<= 32) CONSECUTIVE problems starting at q0 into the warp's columns 0..cnt-1: each
// input array is one contiguous run, read with fully coalesced, independent 8-byte requests (unrolled so that
// many are in flight) and scattered into the transposed layout.
template <int N, int K>
__device__ __forceinline__ void load_block_coop(const QPArgs& a, long long q0, int cnt, double* __restrict__ wcol,
                                                int LD, int lane) {
  using L = TppLayout<N, K>;
  {
    const double* __restrict__ src = a.G + q0 * (N * N);
    const int total = cnt * N * N;
#pragma unroll 8
    for (int it = 0; it < N * N; ++it) {
      const int idx = it * 32 + lane;
      if (idx < total) {
        const int pr = idx / (N * N), e = idx - pr * (N * N);
        const int cc = e / N, rr = e - cc * N;
        const double v = __ldg(src + idx);
        if (rr >= cc) wcol[pr + (L::OG + lidx(rr, cc)) * LD] = v;
      }
    }
  }
  auto run = [&](const double* __restrict__ src, int per, int slot0) {
    const int total = cnt * per;
#pragma unroll 8
    for (int it = 0; it < per; ++it) {
      const int idx = it * 32 + lane;
      if (idx < total) {
        const int pr = idx / per, e = idx - pr * per;
        wcol[pr + (slot0 + e) * LD] = __ldg(src + idx);
      }
    }
  };
  if (K > 0) run(a.A + q0 * (N * K), N * K, L::OA);
  run(a.c + q0 * N, N, L::OC);
  if (K > 0) run(a.b + q0 * K, K, L::OB);
  run(a.ib + q0 * L::M, L::M, L::OIB);
}

// ComputeInitialGuess tail (qp.cc:462-481) for variable v: clamp x into the box in list order
// (ClampX, qp.hpp:43-53), s = max(1e-9, a x + b), z = 1/s.
__device__ __forceinline__ void clamp_and_init_pair(double& xv, double b0, double b1, double& s0, double& s1,
                                                    double& z0, double& z1) {
  xv = xv > -b0 ? xv : -b0;  // a = +1: max(x, -b/a)
  xv = xv < b1 ? xv : b1;    // a = -1: min(x, b/-a)
  s0 = fmax(1.0e-9, xv + b0);
  s1 = fmax(1.0e-9, b1 - xv);
  z0 = 1.0 / s0;
  z1 = 1.0 / s1;
}

// Contiguous run of COUNT doubles owned by one thread: 16-byte vector requests when COUNT is even (the run is
// then 16-byte aligned whenever the array base is; the host checks the bases), scalar otherwise.
template <int COUNT>
__device__ __forceinline__ void load_run(const double* __restrict__ p, double (&out)[COUNT > 0 ? COUNT : 1]) {
  if (COUNT % 2 == 0) {
    const double2* __restrict__ p2 = reinterpret_cast<const double2*>(p);
#pragma unroll
    for (int i = 0; i < COUNT / 2; ++i) {
      const double2 v = __ldg(p2 + i);
      out[2 * i] = v.x;
      out[2 * i + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < COUNT; ++i) out[i] = __ldg(p + i);
  }
}

template <int COUNT>
__device__ __forceinline__ void store_run(double* __restrict__ p, const double (&in)[COUNT]) {
  if (COUNT % 2 == 0) {
    double2* __restrict__ p2 = reinterpret_cast<double2*>(p);
#pragma unroll
    for (int i = 0; i < COUNT / 2; ++i) p2[i] = make_double2(in[2 * i], in[2 * i + 1]);
  } else {
#pragma unroll
    for (int i = 0; i < COUNT; ++i) p[i] = in[i];
  }
}

// ---------------------------------------------------------------------------------------------
// Kernel 1: ComputeInitialGuess (qp.cc:439-482) for every problem -> start state in `vars`.
// NAIVE: x = y = 0. SOLVE_EQUALITY_CONSTRAINED: the KKT system without Sigma at x = y = 0 (r_d = c,
// r_pe = b_eq), factored and solved in registers. A failed factorization is recorded in status[] (= 2) and
// the iterate kernel skips the problem; otherwise status[] = -1 (in flight).
// One thread per problem, no shared memory: each thread streams its own contiguous slices (every sector it
// touches is fully used) with independent vector loads, so the kernel runs at memory speed.
// ---------------------------------------------------------------------------------------------
template <int N, int K>
__global__ void __launch_bounds__(128) qp_tpp_guess_kernel(const QPArgs a) {
  using L = TppLayout<N, K>;
  constexpr int P = L::P, M = L::M, KD = K > 0 ? K : 1, NKD = N * K > 0 ? N * K : 1;
  const bool eq_guess = a.guess == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED;
  for (long long prob = (long long)blockIdx.x * blockDim.x + threadIdx.x; prob < a.batch;
       prob += (long long)gridDim.x * blockDim.x) {
    double sol[P], ib[M];
    load_run<M>(a.ib + prob * M, ib);
    bool ok = true;
    if (eq_guess) {
      double Km[L::NK], Gf[N * N], Af[NKD], cf[N], bf[KD];
      load_run<N * N>(a.G + prob * (N * N), Gf);
      load_run<N * K>(a.A + prob * (N * K), Af);
      load_run<N>(a.c + prob * N, cf);
      load_run<K>(a.b + prob * K, bf);
#pragma unroll
      for (int i = 0; i < N; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) Km[lidx(i, j)] = Gf[j * N + i];  // column-major (i, j)
        sol[i] = -cf[i];
      }
#pragma unroll
      for (int r = 0; r < K; ++r) {
#pragma unroll
        for (int j = 0; j < N; ++j) Km[lidx(N + r, j)] = Af[j * K + r];
#pragma unroll
        for (int q = 0; q <= r; ++q) Km[lidx(N + r, N + q)] = 0.0;
        sol[N + r] = -bf[r];
      }
      ok = ldlt_solve_inreg<P>(Km, sol);  // sol = [x; -y]
    } else {
#pragma unroll
      for (int i = 0; i < P; ++i) sol[i] = 0.0;
    }
    double out[L::STRIDE];
#pragma unroll
    for (int v = 0; v < N; ++v) {
      double xv = ok ? sol[v] : 0.0, s0, s1, z0, z1;
      clamp_and_init_pair(xv, ib[2 * v], ib[2 * v + 1], s0, s1, z0, z1);
      out[v] = ok ? xv : 0.0;
      out[N + 2 * v] = ok ? s0 : 0.0;
      out[N + 2 * v + 1] = ok ? s1 : 0.0;
      out[N + M + K + 2 * v] = ok ? z0 : 0.0;
      out[N + M + K + 2 * v + 1] = ok ? z1 : 0.0;
    }
#pragma unroll
    for (int r = 0; r < K; ++r) out[N + M + r] = ok ? -sol[N + r] : 0.0;
    store_run<L::STRIDE>(a.vars + prob * L::STRIDE, out);
    a.status[prob] = ok ? -1 : MO_QP_FAILED_FACTORIZATION;
    if (!ok) {
      a.iters[prob] = 0;
      if (a.kkt_final) { double* kf = a.kkt_final + prob * 4; kf[0] = 0.0; kf[1] = 0.0; kf[2] = 0.0; kf[3] = 0.0; }
      if (a.lagrange) { a.lagrange[prob * 2] = K > 0 ? 0.0 : qnan(); a.lagrange[prob * 2 + 1] = K > 0 ? 0.0 : qnan(); }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Kernel 2: the Newton loop (qp.cc:110-149) from the start state in `vars`.
// ---------------------------------------------------------------------------------------------
// ASYNC = true: the refill never waits for global memory. The inputs and the start state of a new problem are copied
// with 8-byte cp.async requests (LDGSTS) straight into the idle lane's shared-memory column -- x, y into one of
// XY_SLOTS staging rows of the warp -- while the other lanes compute; the lane joins the NEXT round (one idle round
// per problem, ~9 % of the lane rounds at 11 Newton steps per QP) after a cp.async.wait_group that has long been
// satisfied. ASYNC = false is the round-1 refill (LDG into registers, then STS: a global round trip in nearly every
// round, 37 % of the samples of profiles/r02_tpp_n8_ncu_full_summary.txt), kept for the A/B (mo_debug_set_tpp_async).
constexpr int kTppXYSlots = 6;  // new problems per warp per round; what the 2 x 114 KB CTAs leave free on an SM

__device__ __forceinline__ void tpp_cp8(double* smem_dst, const double* gsrc) {
  const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}

// predicated form on a 32-bit shared address: no branch around the copy
__device__ __forceinline__ void tpp_cp8_if(bool pred, unsigned sdst, const double* gsrc) {
  asm volatile("{ .reg .pred p; setp.ne.b32 p, %2, 0; @p cp.async.ca.shared.global [%0], [%1], 8; }" ::"r"(sdst), "l"(gsrc),
               "r"((int)pred) : "memory");
}

template <int N, int K, int WPB>
constexpr size_t tpp_iterate_smem_bytes(bool async) {
  return sizeof(double) * (TppLayout<N, K>::ETOT * (32 * WPB + 1) + (async ? WPB * kTppXYSlots * (N + K) : 0));
}

template <int N, int K, int WPB, int MINB, bool ASYNC>
__global__ void __launch_bounds__(32 * WPB, MINB) qp_tpp_iterate_kernel(const QPArgs a, const int skip_failed) {
  using L = TppLayout<N, K>;
  constexpr int P = L::P, M = L::M, LD = 32 * WPB + 1, STRIDE = L::STRIDE;
  constexpr int KD = K > 0 ? K : 1;
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int PF_AHEAD = 8;  // problems prefetched into L2 ahead of the refill pointer
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* const smcol = smem + threadIdx.x;
  double* const smwarp = smem + warp * 32;

  // contiguous chunk of the batch owned by this warp
  const long long nwarps = (long long)gridDim.x * WPB, gw = (long long)blockIdx.x * WPB + warp;
  const long long chunk = (a.batch + nwarps - 1) / nwarps;
  long long next = gw * chunk < a.batch ? gw * chunk : a.batch;
  const long long wend = next + chunk < a.batch ? next + chunk : a.batch;
  long long pf_next = next;

  const double tol2 = a.kkt_tol * a.kkt_tol;
  const double tau = 0.995;  // qp.cc:192

  double x[N], y[KD];
  double mu = a.initial_mu, alpha_p = 1.0, alpha_d = 1.0;
  long long prob = -1;
  int iters = 0;
  bool live = false, fin = false, first = false;
  bool loading = false;  // ASYNC: copies of this lane's next problem are in flight
  int st_in = -1, xy_slot = 0;
  double* const xy_stage = smem + L::ETOT * LD + warp * (kTppXYSlots * (N + K));  // ASYNC only
  // ASYNC: what this lane copies for EVERY new problem of its warp is fixed -- element u * 32 + lane of G, of the
  // [A | c | b | bounds] run and of the start state -- so sources (array, per-problem stride, offset) and destinations
  // (byte offset inside the column, or inside the x / y staging row: bit 31 set) are computed once.
  constexpr int GLc = (N * N + 31) / 32, RESTc = N * K + N + K + M, RLc = (RESTc + 31) / 32, VLc = (STRIDE + 31) / 32;
  int g_dst[GLc], v_dst[VLc], w_slot[VLc];
  const double* r_src[RLc];
  int r_stride[RLc];
  const unsigned warp_s = static_cast<unsigned>(__cvta_generic_to_shared(smwarp));
  const unsigned stage_s = static_cast<unsigned>(__cvta_generic_to_shared(xy_stage));
  if constexpr (ASYNC) {
#pragma unroll
    for (int u = 0; u < GLc; ++u) {
      const int e = u * 32 + lane, cc = e / N, rr = e - cc * N;  // column-major (rr, cc)
      g_dst[u] = (e < N * N && rr >= cc) ? (L::OG + lidx(rr, cc)) * LD * 8 : -1;
    }
#pragma unroll
    for (int u = 0; u < RLc; ++u) {
      const int e = u * 32 + lane;
      r_src[u] = nullptr;
      r_stride[u] = 0;
      if (e < RESTc) {
        if (K > 0 && e < N * K) { r_src[u] = a.A + e; r_stride[u] = N * K; }
        else if (e < N * K + N) { r_src[u] = a.c + (e - N * K); r_stride[u] = N; }
        else if (K > 0 && e < N * K + N + K) { r_src[u] = a.b + (e - N * K - N); r_stride[u] = K; }
        else { r_src[u] = a.ib + (e - N * K - N - K); r_stride[u] = M; }
      }
    }
#pragma unroll
    for (int u = 0; u < VLc; ++u) {
      const int e = u * 32 + lane;
      int d = -1, w = -1;
      if (e < N) { d = (int)(0x80000000u | (unsigned)(e * 8)); w = e; }
      else if (e < N + M) { d = (L::OS + (e - N)) * LD * 8; w = L::OS + (e - N); }
      else if (e < N + M + K) { d = (int)(0x80000000u | (unsigned)((N + e - N - M) * 8)); w = N + (e - N - M); }
      else if (e < STRIDE) { d = (L::OZ + (e - N - M - K)) * LD * 8; w = L::OZ + (e - N - M - K); }
      v_dst[u] = e < STRIDE ? d : 0;  // 0 is never a destination of the state (slot OG is G's)
      w_slot[u] = e < STRIDE ? w * LD : -1;
    }
  }
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = 0.0;
#pragma unroll
  for (int i = 0; i < KD; ++i) y[i] = 0.0;

  // (1) as a closure: the cp.async schedule runs it in the MIDDLE of a round (right after the termination test), the register
  // schedule at the top of a round
  auto writeout_and_refill = [&]() {
    // ======== (1) write out finished problems, refill idle lanes ========
    unsigned fmask = __ballot_sync(FULL, fin);
    while (fmask) {  // [x|s|y|z] of lane `src`: x, y parked in slots [0, N+K) of its column, s, z in place
      const int src = __ffs(fmask) - 1;
      fmask &= fmask - 1;
      const long long q = __shfl_sync(FULL, prob, src);
      double* __restrict__ vp = a.vars + q * STRIDE;
      const double* __restrict__ col = smwarp + src;
      if constexpr (ASYNC) {
#pragma unroll
        for (int u = 0; u < VLc; ++u)
          if (w_slot[u] >= 0) vp[u * 32 + lane] = col[w_slot[u]];
      } else {
#pragma unroll
        for (int e0 = 0; e0 < STRIDE; e0 += 32) {
          const int e = e0 + lane;
          if (e < STRIDE) {
            const int slot = e < N ? e : (e < N + M ? L::OS + (e - N) : (e < N + M + K ? N + (e - N - M) : L::OZ + (e - N - M - K)));
            vp[e] = col[slot * LD];
          }
        }
      }
    }
    fin = false;
    const unsigned idle = __ballot_sync(FULL, !live);
    if (ASYNC && idle != 0u && next < wend) {
      __syncwarp();  // the write-out above has read the columns that are about to be overwritten
      const int rank = __popc(idle & ((1u << lane) - 1u));
      const long long left = wend - next;
      int cnt = __popc(idle);
      cnt = cnt < kTppXYSlots ? cnt : kTppXYSlots;
      cnt = (long long)cnt < left ? cnt : (int)left;
      const bool take = !live && rank < cnt;
      if (take) {
        prob = next + rank;
        xy_slot = rank;
        st_in = skip_failed ? a.status[prob] : -1;
      }
      unsigned todo = __ballot_sync(FULL, take);
      long long q = next;  // the taking lanes, in lane order, get consecutive problems and staging rows
      next += cnt;
      for (int sl = 0; todo; ++sl, ++q) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const unsigned col_s = warp_s + src * 8, row_s = stage_s + sl * ((N + K) * 8);
        const double* __restrict__ Gp = a.G + q * (N * N) + lane;
#pragma unroll
        for (int u = 0; u < GLc; ++u) tpp_cp8_if(g_dst[u] >= 0, col_s + g_dst[u], Gp + u * 32);
#pragma unroll
        for (int u = 0; u < RLc; ++u)  // OA, OC, OB, OIB are consecutive slots
          tpp_cp8_if(r_src[u] != nullptr, col_s + (L::OA + u * 32 + lane) * (LD * 8), r_src[u] + q * r_stride[u]);
        const double* __restrict__ vp = a.vars + q * STRIDE + lane;  // start state [x|s|y|z]
#pragma unroll
        for (int u = 0; u < VLc; ++u)
          tpp_cp8_if(v_dst[u] != 0, v_dst[u] < 0 ? row_s + (v_dst[u] & 0x7fffffff) : col_s + v_dst[u], vp + u * 32);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      loading = take;
    }
    if (!ASYNC && idle != 0u && next < wend) {
      const int rank = __popc(idle & ((1u << lane) - 1u));
      const bool take = !live && (next + rank < wend);
      if (take) prob = next + rank;
      const unsigned tmask = __ballot_sync(FULL, take);
      next += __popc(idle);
      if (next > wend) next = wend;
      // the lane's own x, y and the guess kernel's verdict first: their latency overlaps the cooperative loads
      double xin[N], yin[KD];
      if (take) {
        const double* __restrict__ vp = a.vars + prob * STRIDE;
#pragma unroll
        for (int i = 0; i < N; ++i) xin[i] = __ldg(vp + i);
#pragma unroll
        for (int r = 0; r < K; ++r) yin[r] = __ldg(vp + N + M + r);
        if (skip_failed) st_in = a.status[prob];
      }
      // cooperative loads, up to 4 problems per pass: every load of a pass is issued before its first store
      constexpr int GL = (N * N + 31) / 32, REST = N * K + N + K + M, RL = (REST + 31) / 32;
      unsigned todo = tmask;
      while (todo) {
        int src[4];
        long long q[4];
        bool valid[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          valid[g] = todo != 0u;
          src[g] = valid[g] ? __ffs(todo) - 1 : 0;
          todo &= todo - 1;  // 0 stays 0
          q[g] = __shfl_sync(FULL, prob, src[g]);
        }
        double gv[4][GL], rv[4][RL], sz[4][2];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if (valid[g]) {
            const double* __restrict__ Gp = a.G + q[g] * (N * N);
#pragma unroll
            for (int u = 0; u < GL; ++u) {
              const int e = u * 32 + lane;
              gv[g][u] = (e < N * N) ? __ldg(Gp + e) : 0.0;
            }
#pragma unroll
            for (int u = 0; u < RL; ++u) {
              const int e = u * 32 + lane;
              double v = 0.0;
              if (e < REST) {
                if (K > 0 && e < N * K) v = __ldg(a.A + q[g] * (N * K) + e);
                else if (e < N * K + N) v = __ldg(a.c + q[g] * N + (e - N * K));
                else if (K > 0 && e < N * K + N + K) v = __ldg(a.b + q[g] * K + (e - N * K - N));
                else v = __ldg(a.ib + q[g] * M + (e - N * K - N - K));
              }
              rv[g][u] = v;
            }
            const double* __restrict__ vp = a.vars + q[g] * STRIDE;  // start state s, z
            sz[g][0] = (lane < M) ? __ldg(vp + N + lane) : 0.0;
            sz[g][1] = (lane < M) ? __ldg(vp + N + M + K + lane) : 0.0;
          }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if (valid[g]) {
            double* __restrict__ col = smwarp + src[g];
#pragma unroll
            for (int u = 0; u < GL; ++u) {
              const int e = u * 32 + lane;
              const int cc = e / N, rr = e - cc * N;  // column-major (rr, cc)
              if (e < N * N && rr >= cc) col[(L::OG + lidx(rr, cc)) * LD] = gv[g][u];
            }
#pragma unroll
            for (int u = 0; u < RL; ++u) {
              const int e = u * 32 + lane;
              if (e < REST) col[(L::OA + e) * LD] = rv[g][u];  // OA, OC, OB, OIB are consecutive
            }
            if (lane < M) {
              col[(L::OS + lane) * LD] = sz[g][0];
              col[(L::OZ + lane) * LD] = sz[g][1];
            }
          }
        }
      }
      if (take) {
#pragma unroll
        for (int i = 0; i < N; ++i) x[i] = xin[i];
#pragma unroll
        for (int r = 0; r < K; ++r) y[r] = yin[r];
        live = st_in != MO_QP_FAILED_FACTORIZATION;
        first = true;
        iters = 0;
        alpha_p = 1.0; alpha_d = 1.0;
      }
      // L2 prefetch of the lines the next refills will touch
      const long long pf_end = next + PF_AHEAD < wend ? next + PF_AHEAD : wend;
      if (pf_next < next) pf_next = next;
      for (; pf_next < pf_end; ++pf_next) {
        const char* p = nullptr;
        if (lane < 4) p = (const char*)(a.G + pf_next * (N * N)) + lane * 128;
        else if (lane == 4) p = (const char*)(a.ib + pf_next * M);
        else if (lane == 5) p = (const char*)(a.c + pf_next * N);
        else if (lane == 6 && K > 0) p = (const char*)(a.A + pf_next * (N * K));
        else if (lane == 7 && K > 0) p = (const char*)(a.b + pf_next * K);
        else if (lane >= 8 && lane < 11) p = (const char*)(a.vars + pf_next * STRIDE) + (lane - 8) * 128;
        if (p != nullptr && (lane >= 4 || lane * 128 < N * N * 8) && (lane < 8 || (lane - 8) * 128 < STRIDE * 8))
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
      }
      __syncwarp();
    }
  };
  // (5) as a closure: this lane's problem ended -- park x, y for the warp's coalesced store, scalar outputs
  auto park_finished = [&](const int done_status) {
      // ======== (5) this lane's problem ended: park x, y for the warp's coalesced store, scalar outputs ========
#pragma unroll
      for (int i = 0; i < N; ++i) smcol[i * LD] = x[i];
#pragma unroll
      for (int r = 0; r < K; ++r) smcol[(N + r) * LD] = y[r];
      a.status[prob] = done_status;
      a.iters[prob] = iters;
      if (done_status >= MO_QP_FAILED_FACTORIZATION && a.kkt_final) {
        double* __restrict__ kf = a.kkt_final + prob * 4;
        kf[0] = 0.0; kf[1] = 0.0; kf[2] = 0.0; kf[3] = 0.0;
      }
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
      fin = true;
      live = false;
  };

  for (;;) {
    if constexpr (ASYNC) {
      // ======== (0) lanes whose copies were issued one round ago join ========
      if (__any_sync(FULL, loading)) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        if (loading) {
          const double* __restrict__ st = xy_stage + xy_slot * (N + K);
#pragma unroll
          for (int i = 0; i < N; ++i) x[i] = st[i];
#pragma unroll
          for (int r = 0; r < K; ++r) y[r] = st[N + r];
          live = st_in != MO_QP_FAILED_FACTORIZATION;
          first = true;
          iters = 0;
          alpha_p = 1.0; alpha_d = 1.0;
          loading = false;
        }
        __syncwarp();  // the staging rows are free again
      }
    }
    if constexpr (!ASYNC) {
      writeout_and_refill();
      if (!__any_sync(FULL, live)) {
        if (next >= wend && !__any_sync(FULL, loading)) break;
        continue;
      }
    }
    int done_status = -1;
    double rhs[P], hd[N];  // right-hand side, diagonal of G + Sigma

    if (live) {
      // ======== (2) EvaluateKKTConditions (qp.cc:391-420) fused with the assembly of ComputeLDLT (:289-297) ========
      double base[N];  // G_sym x + c - A'y
#pragma unroll
      for (int i = 0; i < N; ++i) base[i] = smcol[(L::OC + i) * LD];
#pragma unroll
      for (int i = 0; i < N; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) {
          const double g = smcol[(L::OG + lidx(i, j)) * LD];
          if (j == i) hd[i] = g;
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
          base[j] = fma(-av, y[r], base[j]);
          pe = fma(av, x[j], pe);
        }
        rhs[N + r] = -pe;
        npe2 = fma(pe, pe, npe2);
      }
      double tv[N];
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
        hd[v] = (hd[v] + w0) + w1;
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
        base[v] += fma(w0, p0, -(w1 * p1));
        tv[v] = i0 - i1;
      }
      // ---- Solve loop bookkeeping (qp.cc:110-146): errors of the step just taken, termination, mu ----
      const double mu_c = sum / (double)M;  // ComputeMu, qp.cc:509-516
      if (first) {
        first = false;
        mu = a.init_mu_with_comp ? mu_c : a.initial_mu;  // qp.cc:115
      } else {
        const double comp2 = comp_err2(sq, sum, mu, M);  // kkt_final uses the mu the step was taken with
        const double max2 = fmax(fmax(nd2, comp2), fmax(npe2, npi2));
        if (a.iter_log) {
          double* __restrict__ row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
          row[4] = sqrt(nd2); row[5] = sqrt(comp2); row[6] = sqrt(npe2); row[7] = sqrt(npi2);
          row[8] = mu; row[9] = alpha_p; row[10] = alpha_d;
          row[11] = qnan(); row[12] = qnan(); row[13] = qnan();
        }
        ++iters;
        if (max2 < tol2 && mu_c < a.comp_tol) {  // qp.cc:132-137
          done_status = MO_QP_SATISFIED_KKT_TOL;
        } else if (iters >= a.max_iterations) {
          done_status = MO_QP_MAX_ITERATIONS;
        } else if (max2 <= mu * mu || !a.decrease_mu_only_on_small_error) {  // qp.cc:140-146
          mu = (a.strategy == MO_BARRIER_FIXED_DECREASE) ? mu * a.sigma : a.sigma * mu_c;
        }
        if (done_status >= 0 && a.kkt_final) {
          double* __restrict__ kf = a.kkt_final + prob * 4;
          kf[0] = sqrt(nd2); kf[1] = sqrt(comp2); kf[2] = sqrt(npe2); kf[3] = sqrt(npi2);
        }
      }
      if (done_status < 0 && !spos) done_status = MO_QP_NONPOSITIVE_SLACK;
      if (done_status < 0 && a.iter_log) {  // kkt_initial of the coming iteration (ComputeErrors with the new mu, qp.cc:118)
        double* __restrict__ row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
        row[0] = sqrt(nd2); row[1] = sqrt(comp_err2(sq, sum, mu, M)); row[2] = sqrt(npe2); row[3] = sqrt(npi2);
      }
#pragma unroll
      for (int v = 0; v < N; ++v) rhs[v] = -fma(-mu, tv[v], base[v]);
    }

    if constexpr (ASYNC) {
      // The termination test sits at the top of a round, so a lane that just ended is known HERE, with three quarters of
      // the round (LDL', solve, update) still ahead: its result is written out and the copies of its next problem are
      // issued now, they land under the rest of this round and the lane joins at the top of the next one -- a finished
      // lane costs no extra idle round (it cost one when (1) ran at the top of the following round).
      if (live && done_status >= 0) park_finished(done_status);
      __syncwarp();  // parked x, y are visible to the lanes that write them out
      writeout_and_refill();
      if (!__any_sync(FULL, live)) {
        if (next >= wend && !__any_sync(FULL, loading)) break;
        continue;
      }
    }
    // ======== (3) reduced KKT matrix [[G + Sigma, .],[A, 0]] (qp.cc:289-297) into registers, LDL', solve:
    //              rhs <- [dx; -dy]. Only the diagonal differs from the copy in shared memory. ========
    double Km[L::NK];
#pragma unroll
    for (int i = 0; i < N; ++i) {
#pragma unroll
      for (int j = 0; j < i; ++j) Km[lidx(i, j)] = smcol[(L::OG + lidx(i, j)) * LD];
      Km[lidx(i, i)] = hd[i];
    }
#pragma unroll
    for (int r = 0; r < K; ++r) {
#pragma unroll
      for (int j = 0; j < N; ++j) Km[lidx(N + r, j)] = smcol[(L::OA + j * K + r) * LD];
#pragma unroll
      for (int q = 0; q <= r; ++q) Km[lidx(N + r, N + q)] = 0.0;
    }
    const bool ok = ldlt_solve_inreg<P>(Km, rhs);
    if (live && done_status < 0 && !ok) done_status = MO_QP_FAILED_FACTORIZATION;  // qp.cc:303-307, per problem

    if (live && done_status < 0) {
      // ======== (4) ds, dz (qp.cc:359-363), step lengths (:485-507), state update (:196-199) ========
      double ds[M], dz[M];
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
        smcol[(L::OS + i) * LD] = fma(ds[i], alpha_p, smcol[(L::OS + i) * LD]);
        smcol[(L::OZ + i) * LD] = fma(dz[i], alpha_d, smcol[(L::OZ + i) * LD]);
      }
    }

    if (live && done_status >= 0) park_finished(done_status);  // cp.async schedule: only a failed factorization gets here
    __syncwarp();
  }
}

}  // namespace mo
