// "warp16": warp-per-problem interior point QP solve for every SMALL shape (n + k <= 16, m <= 32) that has no
// thread-per-problem instantiation: general (variable, a, b) constraint lists, ragged per-problem constraint counts
// (the reference's own GenerateRandomQP batches, test/qp_test.cc:483-524), n = 6 / 10 / 12 ..., all barrier strategies
// and initial-guess methods behind the same entry point (mo_qp_ip_solve_batch).
//
// Replaces, per problem: QPInteriorPointSolver::Setup + Solve (source/qp.cc:20-73, 100-151) and everything below it.
//
// Mapping. One warp owns one problem at a time. The problem's arrays -- G, c, A_eq, b_eq, constraint (var, a, b) lists
// -- are contiguous slices of the batch arrays and are staged into shared memory with ONE 1-D bulk asynchronous copy
// (TMA, cp.async.bulk + mbarrier) per array, double-buffered: the copies of the warp's NEXT problem are in flight while
// the current one iterates (tma.cuh; slices that are not 16-byte multiples are loaded by the lanes instead). The solve
// itself is wip_solve (qp_warp.cuh): the whole primal-dual state in registers, lane L = KKT row L, unpivoted LDL' with
// shuffles, one factorization per Newton step. Problems outside its envelope (PREDICTOR_CORRECTOR, USER_PROVIDED
// start, more than two constraints on one variable, an iteration log requested) run through gen_solve<32> -- the
// Eigen-pivot-order solver of the generic kernel -- on the same staged data, in the same launch.
// Problems are handed out through a ticket counter (iteration counts differ 3..30 per problem).
#pragma once
#include "qp_warp.cuh"
#include "tma.cuh"

namespace mo {

struct W16Stage {  // byte offsets inside one staging buffer
  size_t G, c, A, b, ia, ib, ivar, total;
};

__host__ __device__ inline W16Stage w16_stage_layout(int n, int k, int m, bool box) {
  W16Stage s{};
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t o = off; off = align16(off + bytes); return o; };
  s.G = take(sizeof(double) * n * n);
  s.c = take(sizeof(double) * n);
  s.A = take(sizeof(double) * n * k);
  s.b = take(sizeof(double) * k);
  s.ib = take(sizeof(double) * m);
  s.ia = take(box ? 0 : sizeof(double) * m);
  s.ivar = take(box ? 0 : sizeof(int) * m);
  s.total = off;
  return s;
}

__host__ __device__ inline size_t w16_smem_bytes(int n, int k, int m, bool box) {
  // [2 mbarriers (16 B)] [stage 0] [stage 1] [gen workspace]
  return 16 + 2 * w16_stage_layout(n, k, m, box).total + sizeof(double) * gen_smem_doubles(n, k, m, 32);
}

constexpr int kW16TicketSlots = 64;
__device__ unsigned long long g_w16_tickets[kW16TicketSlots];

// Issue the loads of problem `prob` into staging buffer `buf`; returns nothing: completion through bar (TMA part) and
// program order + __syncwarp (lane part).
__device__ __forceinline__ void w16_issue(const QPArgs& a, long long prob, char* buf, const W16Stage& S, uint64_t* bar,
                                          bool tma, int lane) {
  const int n = a.n, k = a.k, m = a.m;
  const bool box = (a.flags & MO_QP_FLAG_BOX_PAIRS) != 0;
  unsigned tx = 0;
  tx += stage_array(buf + S.G, a.G + prob * (long long)n * n, sizeof(double) * n * n, bar, tma, lane);
  tx += stage_array(buf + S.c, a.c + prob * (long long)n, sizeof(double) * n, bar, tma, lane);
  if (k > 0) {
    tx += stage_array(buf + S.A, a.A + prob * (long long)n * k, sizeof(double) * n * k, bar, tma, lane);
    tx += stage_array(buf + S.b, a.b + prob * (long long)k, sizeof(double) * k, bar, tma, lane);
  }
  if (m > 0) {
    tx += stage_array(buf + S.ib, a.ib + prob * (long long)m, sizeof(double) * m, bar, tma, lane);
    if (!box) {
      tx += stage_array(buf + S.ia, a.ia + prob * (long long)m, sizeof(double) * m, bar, tma, lane);
      tx += stage_array(buf + S.ivar, a.ivar + prob * (long long)m, sizeof(int) * m, bar, tma, lane);
    }
  }
  if (lane == 0) mbar_expect_tx(bar, tx);  // the one pending arrival of this phase (+ the bytes in flight)
}

template <int PMAX>
__global__ void __launch_bounds__(32) qp_warp16_kernel(const QPArgs a, unsigned long long* __restrict__ ticket,
                                                       const int use_tma) {
  extern __shared__ __align__(16) unsigned char w16_smem[];
  const int lane = threadIdx.x;
  const int n = a.n, k = a.k, mmax = a.m;
  const bool box = (a.flags & MO_QP_FLAG_BOX_PAIRS) != 0, tma = use_tma != 0;
  const W16Stage S = w16_stage_layout(n, k, mmax, box);
  uint64_t* bars = reinterpret_cast<uint64_t*>(w16_smem);
  char* stage[2] = {reinterpret_cast<char*>(w16_smem) + 16, reinterpret_cast<char*>(w16_smem) + 16 + S.total};
  double* work = reinterpret_cast<double*>(w16_smem + 16 + 2 * S.total);
  const long long vstride = n + 2LL * mmax + k;
  const GenSolveParams sp{a.initial_mu, a.sigma, a.kkt_tol, a.comp_tol, a.max_iterations, a.strategy,
                          a.decrease_mu_only_on_small_error, a.guess, a.init_mu_with_comp};
  if (lane == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
  __syncwarp();
  unsigned phase[2] = {0u, 0u};
  auto next_ticket = [&]() -> long long {
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1ULL);
    return (long long)__shfl_sync(kFull, t, 0);
  };
  long long prob = next_ticket();
  int cur = 0;
  if (prob < a.batch) w16_issue(a, prob, stage[0], S, &bars[0], tma, lane);
  while (prob < a.batch) {
    // prefetch the next problem into the other buffer (its previous contents were consumed one trip ago)
    const long long nxt = next_ticket();
    if (nxt < a.batch) {
      fence_proxy_async();
      w16_issue(a, nxt, stage[cur ^ 1], S, &bars[cur ^ 1], tma, lane);
    }
    mbar_wait(&bars[cur], phase[cur]);
    phase[cur] ^= 1u;
    __syncwarp();
    const char* buf = stage[cur];
    GenWork W;
    GenProb Q;
    const int m = a.icount ? a.icount[prob] : mmax;
    gen_carve(W, work, n, k, m, 32);
    Q.G = reinterpret_cast<const double*>(buf + S.G);
    Q.c = reinterpret_cast<const double*>(buf + S.c);
    Q.A = k ? reinterpret_cast<const double*>(buf + S.A) : nullptr;
    Q.b = k ? reinterpret_cast<const double*>(buf + S.b) : nullptr;
    {
      const double* ib = reinterpret_cast<const double*>(buf + S.ib);
      const double* ia = reinterpret_cast<const double*>(buf + S.ia);
      const int* iv = reinterpret_cast<const int*>(buf + S.ivar);
      for (int i = lane; i < m; i += 32) {
        W.cv[i] = box ? i / 2 : iv[i];
        W.ca[i] = box ? ((i & 1) ? -1.0 : 1.0) : ia[i];
        W.cb[i] = ib[i];
      }
    }
    if (a.guess == MO_GUESS_USER_PROVIDED) gen_load_vars<32>(W, mmax, a.vars + prob * vstride, W.vars);
    else { for (int i = lane; i < W.T; i += 32) W.vars[i] = 0.0; }
    __syncwarp();
    int iters = 0, status = -1;
    double kkt_after[4];
    if (!a.iter_log) status = wip_solve<PMAX>(W, Q, sp, iters, kkt_after);
    if (status < 0)
      status = gen_solve<32>(W, Q, sp, a.iter_log ? a.iter_log + prob * a.max_iterations * MO_QP_ITER_LOG_STRIDE : nullptr,
                             iters, kkt_after);
    __syncwarp();
    gen_store_vars<32>(W, mmax, W.vars, a.vars + prob * vstride);
    if (lane == 0) {
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
    __syncwarp();
    prob = nxt;
    cur ^= 1;
  }
}

}  // namespace mo
