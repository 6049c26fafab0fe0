// CTA-per-problem interior point QP solve for medium dense box-constrained QPs (n, k multiples of 8, n + k <= 80)
// with the O(p^3) work on the FP64 tensor path: mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4).
//
// Replaces, for every problem of a batch: QPInteriorPointSolver::Setup + Solve (source/qp.cc:20-73, 100-151):
// ComputeInitialGuess (:439-482), the Iterate loop (:153-201), EvaluateKKTConditions (:391-420), ComputeLDLT
// (:275-316) + SolveForUpdate (:318-364) as ONE unpivoted LDL' factorization and ONE two-sided solve (no explicit
// inverse), ComputeAlpha (:485-507), ComputeMu (:509-516), ComputeErrors (:423-437).
//
// Mapping (P = n + k, NT = P/8 tile rows, W warps per CTA):
//   * The reduced KKT matrix [[G + Sigma, A'],[A, 0]] (lower triangle, 8x8 tiles) lives in DMMA ACCUMULATOR
//     fragments: tile row I belongs to warp rowwarp(I) (snake order, balanced as the factorization shrinks).
//   * Right-looking LDL' over 8-column panels. Per panel: owners spill the panel's tiles to shared memory; one warp
//     factors the 8x8 diagonal block (lane = row, shuffles); one THREAD per panel row does the triangular solve and
//     writes L in a 4-column "chunk" layout whose rows are exactly the DMMA A/B fragment order (conflict-free LDS.64);
//     the trailing update is 2 DMMAs per tile, A fragments shared along a tile row.
//   * The right-hand side rides along as an extra row of the panel, so the forward substitution is free; the back
//     substitution runs block by block from L in shared memory.
//   * G and A_eq are re-staged into shared memory once per Newton step with cp.async (they stay L2-resident for the
//     life of the problem) into the region L will overwrite: packed trapezoid columns padded so that both the
//     fragment loads and the thread-per-row symmetric matvec are bank-conflict free.
#pragma once
#include <type_traits>

#include "qp_common.cuh"

namespace mo {

template <int N, int K, int W>
struct DmLayout {
  static constexpr int P = N + K, NT = P / 8, NTN = N / 8, M = 2 * N, THREADS = 32 * W;
  static_assert(N % 8 == 0 && K % 8 == 0 && N > 0, "tile shapes");
  static_assert(P <= THREADS, "one thread per KKT row");
  static_assert(2 * N <= 4 * THREADS, "matvec task split");
  __host__ __device__ static constexpr int rowwarp(int I) {
    const int idx = NT - 1 - I, r = idx % (2 * W);
    return r < W ? r : 2 * W - 1 - r;
  }
  __host__ __device__ static constexpr int slotbase(int I) {
    int s = 0;
    for (int i = 0; i < I; ++i)
      if (rowwarp(i) == rowwarp(I)) s += i + 1;
    return s;
  }
  __host__ __device__ static constexpr int maxslots() {
    int best = 0;
    for (int w = 0; w < W; ++w) {
      int s = 0;
      for (int i = 0; i < NT; ++i)
        if (rowwarp(i) == w) s += i + 1;
      best = s > best ? s : best;
    }
    return best;
  }
  // L in 4-column chunks: chunk kc holds columns 4kc..4kc+3 of rows 8*(kc/2) .. P-1, row-major [row][4], +4 pad
  __host__ __device__ static constexpr int chunk_off(int kc) {
    const int j = kc / 2;
    return 8 * j * (P + 1) - 32 * j * (j - 1) + ((kc & 1) ? (P - 8 * j) * 4 + 4 : 0);
  }
  static constexpr int LSIZE = chunk_off(2 * NT);
  // staged G: column c (rows >= 8*(c/8)) of length N - 8*(c/8) + 2
  __host__ __device__ static constexpr int gcol_off(int c) {
    const int jb = c / 8;
    return 8 * (jb * (N + 2) - 4 * jb * (jb - 1)) + (c % 8) * (N - 8 * jb + 2);
  }
  static constexpr int GSIZE = gcol_off(N);
  static constexpr int ALD = K + 2;
  static constexpr int ASIZE = K > 0 ? N * ALD : 0;
  static constexpr int REGION = (LSIZE > GSIZE + ASIZE ? LSIZE : GSIZE + ASIZE) + 2;
  static constexpr int PLD = 10;  // row stride of the raw panel buffer
  // shared memory map (doubles)
  static constexpr int O_REGION = 0;
  static constexpr int O_PRAW = O_REGION + REGION;
  static constexpr int O_DJ = O_PRAW + P * PLD;       // 2 x {d[8], 1/d[8]}, rhs exchange [8], w (unscaled) [8]
  static constexpr int O_X = O_DJ + 48;
  static constexpr int O_Y = O_X + N;
  static constexpr int O_S = O_Y + (K > 0 ? K : 2);
  static constexpr int O_Z = O_S + M;
  static constexpr int O_C = O_Z + M;
  static constexpr int O_B = O_C + N;
  static constexpr int O_IB = O_B + (K > 0 ? K : 2);
  static constexpr int O_SIG = O_IB + M;              // Sigma_v = z0/s0 + z1/s1
  static constexpr int O_T1 = O_SIG + N;              // matvec partials / base
  static constexpr int O_T2 = O_T1 + N;
  static constexpr int O_RPE = O_T2 + N;
  static constexpr int O_DALL = O_RPE + (K > 0 ? K : 2);  // all pivots d[P]
  static constexpr int O_WP = O_DALL + P;             // forward-solved, D^-1-scaled rhs
  static constexpr int O_SOL = O_WP + P;              // solution [dx; -dy]
  static constexpr int O_RED = O_SOL + P;             // reductions: 8 values x W warps (+ flags)
  static constexpr int O_ONES = O_RED + 8 * W + 8;    // 8 x 1.0 and an 8x8 identity (row stride PLD): let the threads that
  static constexpr int O_EYE = O_ONES + 8;            // invert L_jj run the P3 row recurrence without a divergent path
  static constexpr int TOTAL = O_EYE + 8 * PLD;
  static constexpr size_t SMEM_BYTES = sizeof(double) * TOTAL;
};

#ifdef MO_DM_TRACE
// Debug build only: thread 0 of CTA 0 stamps phase boundaries into a shared-memory ring (cheap), dumped at kernel end.
__device__ long long g_dm_trace[2048];
__device__ int g_dm_trace_n;
__shared__ long long s_dm_trace[1024];
__shared__ int s_dm_trace_n;
#define DM_T(tag) do { if (blockIdx.x == 0 && threadIdx.x == 0 && s_dm_trace_n < 1024) { s_dm_trace[s_dm_trace_n++] = (clock64() << 8) | (tag); } } while (0)
#define DM_T_INIT() do { if (threadIdx.x == 0) s_dm_trace_n = 0; __syncthreads(); } while (0)
#define DM_T_DUMP() do { __syncthreads(); if (blockIdx.x == 0 && threadIdx.x == 0) { for (int i_ = 0; i_ < s_dm_trace_n; ++i_) g_dm_trace[i_] = s_dm_trace[i_]; g_dm_trace_n = s_dm_trace_n; } } while (0)
#else
#define DM_T(tag) do {} while (0)
#define DM_T_INIT() do {} while (0)
#define DM_T_DUMP() do {} while (0)
#endif

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c0), "+d"(c1)
      : "d"(a), "d"(b));
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Reciprocal with a 3-deep dependent FMA chain after MUFU.RCP64H: x(1 + e + e^2), e = 1 - d x, |e| < 2^-20, so the
// truncation error e^3 < 2^-60 is below double rounding; used on the pivot chain of the diagonal block.
__device__ __forceinline__ double rcp_chain(double d) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
  const double e = fma(-d, x, 1.0);
  const double p = fma(e, e, e);
  return fma(x, p, x);
}

// Sum / min over the CTA of NV values per thread; result broadcast to every thread. `red` holds NV*W doubles.
// `between` runs between the two barriers of the reduction (work that needs the first barrier's visibility and whose
// shared-memory reads must be complete, CTA-wide, when the reduction returns).
template <int W, int NV, bool IS_MIN, class F>
__device__ __forceinline__ void block_reduce(double (&v)[NV], double* red, F&& between) {
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double t = __shfl_xor_sync(0xffffffffu, v[i], o);
      v[i] = IS_MIN ? fmin(v[i], t) : v[i] + t;
    }
  if (W == 1) {
    __syncwarp();
    between();
    __syncwarp();
    return;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0)
#pragma unroll
    for (int i = 0; i < NV; ++i) red[warp * NV + i] = v[i];
  between();
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    double t = red[i];
#pragma unroll
    for (int w = 1; w < W; ++w) t = IS_MIN ? fmin(t, red[w * NV + i]) : t + red[w * NV + i];
    v[i] = t;
  }
}
template <int W, int NV, bool IS_MIN>
__device__ __forceinline__ void block_reduce(double (&v)[NV], double* red) {
  block_reduce<W, NV, IS_MIN>(v, red, [] {});
}

// cp.async staging of the problem's G (lower trapezoid) and A_eq into the region L will later overwrite.
template <int N, int K, int W>
__device__ __forceinline__ void dm_stage_issue(double* __restrict__ region, const double* __restrict__ Gp,
                                               const double* __restrict__ Ap) {
  using L = DmLayout<N, K, W>;
  constexpr int HG = N / 2;
  for (int task = threadIdx.x; task < N * HG; task += L::THREADS) {
    const int c = task / HG, row = 2 * (task - c * HG), rb = 8 * (c / 8);
    if (row >= rb) cp_async16(region + L::gcol_off(c) + row - rb, Gp + c * N + row);
  }
  if (K > 0) {
    constexpr int HA = K / 2;
    for (int task = threadIdx.x; task < N * HA; task += L::THREADS) {
      const int c = task / HA, row = 2 * (task - c * HA);
      cp_async16(region + L::GSIZE + c * L::ALD + row, Ap + c * K + row);
    }
  }
  cp_async_commit();
}

// Reduced KKT matrix [[G + Sigma, .],[A, 0]] (ComputeLDLT assembly, qp.cc:289-297) into accumulator fragments.
template <int N, int K, int W, int MAXS>
__device__ __forceinline__ void dm_assemble(double (&acc)[MAXS][2], const double* __restrict__ sm, bool with_sigma, int tid) {
  using L = DmLayout<N, K, W>;
  const int lane = tid & 31, warp = tid >> 5, r = lane >> 2, q = lane & 3;
  const double* __restrict__ region = sm + L::O_REGION;
#pragma unroll
  for (int I = 0; I < L::NT; ++I) {
    if (L::rowwarp(I) != warp) continue;
#pragma unroll
    for (int J = 0; J <= I; ++J) {
      const int slot = L::slotbase(I) + J;
      double c0 = 0.0, c1 = 0.0;
      if (I < L::NTN) {
          const int len = N - 8 * J + 2;
        const int base = L::gcol_off(8 * J) + 8 * (I - J) + r;
        c0 = region[base + (2 * q) * len];
        c1 = region[base + (2 * q + 1) * len];
        if (I == J && with_sigma) {
          const double sg = sm[L::O_SIG + 8 * I + r];
          if (r == 2 * q) c0 += sg;
          if (r == 2 * q + 1) c1 += sg;
        }
      } else if (J < L::NTN) {
        const int base = L::GSIZE + 8 * (I - L::NTN) + r;
        c0 = region[base + (8 * J + 2 * q) * L::ALD];
        c1 = region[base + (8 * J + 2 * q + 1) * L::ALD];
      }
      acc[slot][0] = c0;
      acc[slot][1] = c1;
    }
  }
}

// LDL' of the 8x8 diagonal block of panel j (raw values in praw rows 8j..8j+7): lane = row, pivots taken in PAIRS so
// that the two reciprocals of a pair overlap (1/d1 = a00 / det): 4 dependent reciprocal latencies per block, not 8.
// Writes d, 1/d into djb[0..15], the pivots into DALL and the unit lower L_jj into chunks (offA, offB) rows 0..7.
template <int N, int K, int W>
__device__ __forceinline__ void dm_diag_block(double* __restrict__ sm, int j, double* __restrict__ djb, int offA, int offB) {
  using L = DmLayout<N, K, W>;
  const int lane = threadIdx.x & 31, lr = lane & 7;
  const double* __restrict__ praw = sm + L::O_PRAW;
  double* __restrict__ Lch = sm + L::O_REGION;
  double a[8];
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    const double2 v = *reinterpret_cast<const double2*>(praw + (8 * j + lr) * L::PLD + c);
    a[c] = v.x; a[c + 1] = v.y;
  }
  // Pivot bookkeeping without branches: every lane computes the same pivots; lane c keeps (d_c, 1/d_c) by select and
  // the 8 pairs are stored once at the end. A zero / non-finite pivot turns its reciprocal chain into NaN
  // (rcp.approx(0) = inf, inf * 0 = NaN), which `chk` accumulates: one FMA per pivot instead of four compares.
  double chk = 0.0, myd = 0.0, myinv = 0.0;
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    const double a00 = __shfl_sync(0xffffffffu, a[c], c);
    const double a10 = __shfl_sync(0xffffffffu, a[c], c + 1);
    const double a11 = __shfl_sync(0xffffffffu, a[c + 1], c + 1);
    const double det = fma(a00, a11, -(a10 * a10));
    const double inv0 = rcp_chain(a00), rdet = rcp_chain(det);
    const double inv1 = a00 * rdet;               // 1 / d1 with d1 = det / a00
    const double d1 = fma(-(a10 * inv0), a10, a11);
    chk = fma(inv0, 0.0, chk);
    chk = fma(inv1, 0.0, chk);
    const double u0 = a[c], l0 = u0 * inv0;
    const double u1 = fma(-l0, a10, a[c + 1]), l1 = u1 * inv1;
    // (a 2x2 block-pivot form whose only post-reciprocal op is one FMA was tried: more issue slots, measured slower)
#pragma unroll
    for (int c2 = c + 2; c2 < 8; ++c2) {
      const double v0 = __shfl_sync(0xffffffffu, u0, c2), v1 = __shfl_sync(0xffffffffu, u1, c2);
      a[c2] = fma(-l1, v1, fma(-l0, v0, a[c2]));
    }
    myd = lr == c ? a00 : (lr == c + 1 ? d1 : myd);
    myinv = lr == c ? inv0 : (lr == c + 1 ? inv1 : myinv);
    // unit lower L_jj, columns c, c+1 of row lr (chunk h = c / 4, 16-byte slot (c & 2))
    const double w0 = lr > c ? l0 : (lr == c ? 1.0 : 0.0), w1 = lr > c + 1 ? l1 : (lr == c + 1 ? 1.0 : 0.0);
    if (lane < 8) *reinterpret_cast<double2*>(Lch + (c >= 4 ? offB : offA) + lr * 4 + (c & 2)) = make_double2(w0, w1);
  }
  if (lane < 8) {
    djb[lr] = myd; djb[8 + lr] = myinv;
    sm[L::O_DALL + 8 * j + lr] = myd;
  }
  if (chk != chk && lane == 0) *reinterpret_cast<int*>(sm + L::O_RED + 8 * W) = 1;
}

// Trailing update of panel j = C - 1 with the panel index STATIC (dispatched by dm_switch below): the accumulator slots
// are register names, so a runtime j would force every one of the NT(NT+1)/2 tile bodies to be issued predicated-off on
// every panel (measured: 1000 of the 3400 cycles of a panel). Pass A updates and spills tile column C (raw values for
// P3 / the look-ahead diagonal block), pass B updates the columns right of it.
// f(integral_constant<int, c>) for the runtime c in [0, NT): one jump, static tile code behind it.
template <int NT, class F>
__device__ __forceinline__ void dm_switch(int c, F&& f) {
#define DM_CASE(V) case V: if constexpr (V < NT) f(std::integral_constant<int, V>{}); break;
  switch (c) { DM_CASE(0) DM_CASE(1) DM_CASE(2) DM_CASE(3) DM_CASE(4) DM_CASE(5) DM_CASE(6) DM_CASE(7) DM_CASE(8) DM_CASE(9)
               DM_CASE(10) DM_CASE(11) DM_CASE(12) DM_CASE(13) DM_CASE(14) DM_CASE(15) default: break; }
#undef DM_CASE
  static_assert(NT <= 16, "extend dm_switch");
}

template <int N, int K, int W, int WARP, int C, int MAXS>
__device__ __forceinline__ void dm_pass_a(double (&acc)[MAXS][2], const double* __restrict__ ch0, const double* __restrict__ ch1,
                                          double nd0, double nd1, double* __restrict__ praw, int r, int q) {
  using L = DmLayout<N, K, W>;
  if constexpr (C > 0) {
    const double bn0 = ch0[32 * C], bn1 = ch1[32 * C];  // B fragments of tile column C
    double a0[L::NT], a1[L::NT];
#pragma unroll
    for (int I = C; I < L::NT; ++I)
      if (L::rowwarp(I) == WARP) { a0[I] = ch0[32 * I] * nd0; a1[I] = ch1[32 * I] * nd1; }
    // the two k-halves of a tile are dependent DMMAs: all first halves, then all second halves
#pragma unroll
    for (int I = C; I < L::NT; ++I)
      if (L::rowwarp(I) == WARP) dmma884(acc[L::slotbase(I) + C][0], acc[L::slotbase(I) + C][1], a0[I], bn0);
#pragma unroll
    for (int I = C; I < L::NT; ++I)
      if (L::rowwarp(I) == WARP) dmma884(acc[L::slotbase(I) + C][0], acc[L::slotbase(I) + C][1], a1[I], bn1);
  }
#pragma unroll
  for (int I = C; I < L::NT; ++I)
    if (L::rowwarp(I) == WARP)
      *reinterpret_cast<double2*>(praw + (8 * I + r) * L::PLD + 2 * q) =
          make_double2(acc[L::slotbase(I) + C][0], acc[L::slotbase(I) + C][1]);
}

template <int N, int K, int W, int WARP, int C, int MAXS>
__device__ __forceinline__ void dm_pass_b(double (&acc)[MAXS][2], const double* __restrict__ ch0, const double* __restrict__ ch1,
                                          double nd0, double nd1) {
  using L = DmLayout<N, K, W>;
  if constexpr (C > 0) {
#pragma unroll
    for (int I = C + 1; I < L::NT; ++I) {
      if (L::rowwarp(I) != WARP) continue;
      const double a0 = ch0[32 * I] * nd0, a1 = ch1[32 * I] * nd1;
#pragma unroll
      for (int J2 = C + 1; J2 <= I; ++J2) dmma884(acc[L::slotbase(I) + J2][0], acc[L::slotbase(I) + J2][1], a0, ch0[32 * J2]);
#pragma unroll
      for (int J2 = C + 1; J2 <= I; ++J2) dmma884(acc[L::slotbase(I) + J2][0], acc[L::slotbase(I) + J2][1], a1, ch1[32 * J2]);
    }
  }
}

// f(integral_constant<int, warp>, integral_constant<int, c>) for the runtime (warp, c): one jump, static tile code
// behind it (each tile body exists once: it belongs to one warp and one column).
template <int W, int NT, class F>
__device__ __forceinline__ void dm_switch2(int warp, int c, F&& f) {
  dm_switch<W>(warp, [&](auto ww) { dm_switch<NT>(c, [&](auto cc) { f(ww, cc); }); });
}

// Unpivoted LDL' of the matrix held in `acc` (right-looking, 8-column panels, DMMA trailing updates) fused with the
// forward substitution of the right-hand side `racc` (thread t < P holds entry t), then the back substitution.
// On return sm[O_SOL .. O_SOL+P) = K^-1 rhs for every thread to read. Returns false (uniformly) on a zero or
// non-finite pivot (-> MO_QP_FAILED_FACTORIZATION, the batched stand-in for the exception of qp.cc:303-307).
//
// Schedule per panel j (two CTA barriers): [P3: one thread per row below the block solves against L_jj D_j]  |barrier|
// [P4: every warp first updates and spills its tiles of column j+1; the owner of tile row j+1 then factors the NEXT
// diagonal block (look-ahead) while the other warps finish the trailing update]  |barrier|. The panel loop is a RUNTIME
// loop (tile loops stay unrolled so that the accumulators keep static register names): fully unrolled it is 0.5 MB of
// SASS. j = -1 is the prologue (spill column 0, factor block 0).
template <int N, int K, int W, int MAXS>
__device__ __forceinline__ bool dm_factor_solve(double (&acc)[MAXS][2], double racc, double* __restrict__ sm, int tid) {
  using L = DmLayout<N, K, W>;
  constexpr int P = L::P, NT = L::NT, PLD = L::PLD;
  const int lane = tid & 31, warp = tid >> 5, r = lane >> 2, q = lane & 3;
  double* __restrict__ Lch = sm + L::O_REGION;
  double* __restrict__ praw = sm + L::O_PRAW;
  double* __restrict__ dj = sm + L::O_DJ;  // [0..31] d, 1/d of the panel (double buffered), [32..39] rhs block, [40..47] w
  int* failflag = reinterpret_cast<int*>(sm + L::O_RED + 8 * W);
  double l[8];
#pragma unroll 1
  for (int j = -1; j < NT; ++j) {
    const int off0 = j >= 0 ? L::chunk_off(2 * j) : 0, off1 = off0 + (P - 8 * j) * 4 + 4;
    const double* __restrict__ djc = dj + (j & 1) * 16;
    const bool is_row = tid >= 8 * (j + 1) && tid < P;
    if (j >= 0) {
      // ---- P3: triangular solve, one thread per row below the block; the last thread carries the right-hand side.
      //      The 8 threads of block j itself run the same recurrence on a unit vector: thread 8j + c gets column c of
      //      inv(L_jj) (for the back substitution) at no extra issue cost, parked in its dead row of praw. ----
      const bool is_rhs = tid == L::THREADS - 1, is_inv = tid >= 8 * j && tid < 8 * j + 8;
      if (is_row || is_inv || is_rhs) {  // one instruction stream for the three roles: only the pointers differ
        const double* __restrict__ src = is_rhs ? dj + 32 : (is_inv ? sm + L::O_EYE + (tid & 7) * PLD : praw + tid * PLD);
        const double* __restrict__ sc = is_inv ? sm + L::O_ONES : djc + 8;
        double* __restrict__ d0 = is_rhs ? sm + L::O_WP + 8 * j : (is_inv ? praw + tid * PLD : Lch + off0 + (tid - 8 * j) * 4);
        double* __restrict__ d1 = is_row ? Lch + off1 + (tid - 8 * j) * 4 : d0 + 4;
        double a[8], u[8];
#pragma unroll
        for (int c = 0; c < 8; c += 2) {
          const double2 v = *reinterpret_cast<const double2*>(src + c);
          a[c] = v.x; a[c + 1] = v.y;
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          double v = a[c];
#pragma unroll
          for (int k2 = 0; k2 < c; ++k2) v = fma(-u[k2], Lch[(k2 < 4 ? off0 : off1) + c * 4 + (k2 & 3)], v);
          u[c] = v;
          l[c] = v * sc[c];
        }
        *reinterpret_cast<double2*>(d0) = make_double2(l[0], l[1]);
        *reinterpret_cast<double2*>(d0 + 2) = make_double2(l[2], l[3]);
        *reinterpret_cast<double2*>(d1) = make_double2(l[4], l[5]);
        *reinterpret_cast<double2*>(d1 + 2) = make_double2(l[6], l[7]);
        if (is_rhs) {
#pragma unroll
          for (int c = 0; c < 8; c += 2) *reinterpret_cast<double2*>(dj + 40 + c) = make_double2(u[c], u[c + 1]);
        }
      }
      DM_T(3);
      __syncthreads();
      DM_T(4);
      if (j == NT - 1) break;
    }
    // ---- P4: trailing update C(I,J') -= L(I,j) D_j L(J',j)' with DMMA; column j+1 first, then look-ahead ----
    double nd0 = 0.0, nd1 = 0.0;
    const double* __restrict__ ch0 = Lch + off0 + r * 4 + q - 32 * j;  // + 32*I -> row 8I + r of chunk 2j
    const double* __restrict__ ch1 = Lch + off1 + r * 4 + q - 32 * j;
    if (j >= 0) {
      nd0 = -djc[q]; nd1 = -djc[4 + q];
      if (is_row) {  // forward substitution of the right-hand side rows below the block
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int c = 0; c < 8; c += 2) { s0 = fma(l[c], dj[40 + c], s0); s1 = fma(l[c + 1], dj[41 + c], s1); }
        racc -= s0 + s1;
      }
    }
    if (tid >= 8 * (j + 1) && tid < 8 * (j + 1) + 8) dj[32 + tid - 8 * (j + 1)] = racc;
    // pass A: tiles of column j+1 -> update, spill raw values for the next panel
    dm_switch2<W, NT>(warp, j + 1, [&](auto ww, auto cc) {
      dm_pass_a<N, K, W, decltype(ww)::value, decltype(cc)::value>(acc, ch0, ch1, nd0, nd1, praw, r, q);
    });
    DM_T(8);
    if (warp == L::rowwarp(j + 1)) {  // look-ahead: factor diagonal block j+1 now
      __syncwarp();
      const int offA = L::chunk_off(2 * j + 2);
      dm_diag_block<N, K, W>(sm, j + 1, dj + ((j + 1) & 1) * 16, offA, offA + (P - 8 * (j + 1)) * 4 + 4);
      DM_T(9);
    }
    // pass B: the rest of the trailing matrix
    dm_switch2<W, NT>(warp, j + 1, [&](auto ww, auto cc) {
      dm_pass_b<N, K, W, decltype(ww)::value, decltype(cc)::value>(acc, ch0, ch1, nd0, nd1);
    });
    DM_T(1);
    __syncthreads();
    DM_T(2);
  }
  DM_T(6);
  // The factor is consumed: end the live range of the accumulators here (warps that own fewer tiles never rewrite
  // some slots, so without this the compiler keeps all of them live through the residual / update phases).
#pragma unroll
  for (int s = 0; s < MAXS; ++s) { acc[s][0] = 0.0; acc[s][1] = 0.0; }
  // ---- back substitution L' sol = w' (w' = D^-1 L^-1 rhs from the fused forward pass, in WP) by WARP 0 ALONE: every
  //      block step is a short dependent chain, so CTA barriers between the steps cost more than the parallelism they
  //      buy. Block J is finalised as x_J = inv(L_JJ)' b_J (columns of inv(L_JJ) parked in praw by P3), then rows
  //      t < 8J (lane l carries t = l, l + 32, ...) are updated. A runtime loop: the code stays small. ----
  if (warp == 0) {
    double* __restrict__ WP = sm + L::O_WP;
    double* __restrict__ SOLv = sm + L::O_SOL;
    const int c8 = lane & 7;
#pragma unroll 1
    for (int J = NT - 1; J >= 0; --J) {
      const double* __restrict__ tcp = praw + (8 * J + c8) * PLD;
      const double* __restrict__ bj = WP + 8 * J;
      double x0 = 0.0, x1 = 0.0;
#pragma unroll
      for (int rr = 0; rr < 8; rr += 2) {
        const double2 t2 = *reinterpret_cast<const double2*>(tcp + rr);
        const double2 b2 = *reinterpret_cast<const double2*>(bj + rr);
        x0 = fma(t2.x, b2.x, x0);
        x1 = fma(t2.y, b2.y, x1);
      }
      if (lane < 8) SOLv[8 * J + lane] = x0 + x1;
      __syncwarp();
      if (J == 0) break;
      double sol[8];
#pragma unroll
      for (int rr = 0; rr < 8; rr += 2) {
        const double2 v = *reinterpret_cast<const double2*>(SOLv + 8 * J + rr);
        sol[rr] = v.x; sol[rr + 1] = v.y;
      }
      for (int t = lane; t < 8 * J; t += 32) {
        const double* __restrict__ lp = Lch + L::chunk_off(t >> 2) + (t & 3) - 32 * (t >> 3) + 32 * J;  // L[8J + rr][t]
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int rr = 0; rr < 8; rr += 2) {
          s0 = fma(lp[4 * rr], sol[rr], s0);
          s1 = fma(lp[4 * rr + 4], sol[rr + 1], s1);
        }
        WP[t] -= s0 + s1;
      }
      __syncwarp();
    }
  }
  __syncthreads();
  DM_T(7);
  return *failflag == 0;
}

// ComputeInitialGuess tail for variable v (qp.cc:462-481): clamp in list order (ClampX, qp.hpp:43-53),
// s = max(1e-9, a x + b), z = 1/s.
__device__ __forceinline__ void dm_clamp_init(double* X, double* S, double* Z, const double* IB, int v) {
  double xv = X[v];
  const double b0 = IB[2 * v], b1 = IB[2 * v + 1];
  xv = xv > -b0 ? xv : -b0;
  xv = xv < b1 ? xv : b1;
  const double s0 = fmax(1.0e-9, xv + b0), s1 = fmax(1.0e-9, b1 - xv);
  X[v] = xv;
  S[2 * v] = s0; S[2 * v + 1] = s1;
  Z[2 * v] = 1.0 / s0; Z[2 * v + 1] = 1.0 / s1;
}

// ---------------------------------------------------------------------------------------------------------
// The solve kernel: persistent CTAs, one problem per CTA at a time.
// ---------------------------------------------------------------------------------------------------------
template <int N, int K, int W, int MINB>
__global__ void __launch_bounds__(32 * W, MINB) qp_dm_kernel(const QPArgs a) {
  using L = DmLayout<N, K, W>;
  static_assert(L::P < L::THREADS, "the last thread carries the right-hand side");
  constexpr int P = L::P, M = L::M, THREADS = L::THREADS, MAXS = L::maxslots(), STRIDE = N + 2 * M + K;
  extern __shared__ __align__(16) double sm[];
  // Virtual thread index: the warp roles (tile rows, per-variable work, back substitution) are rotated by the CTA's
  // residency slot so that co-resident CTAs do not pile their heavy warps onto the same SM sub-partition.
  const int rot = W > 1 ? (blockIdx.x / (gridDim.x >= MINB ? gridDim.x / MINB : 1)) & (W - 1) : 0;
  const int tid = ((((int)threadIdx.x >> 5) + rot) & (W - 1)) * 32 + ((int)threadIdx.x & 31);
  double* const X = sm + L::O_X;
  double* const Y = sm + L::O_Y;
  double* const S = sm + L::O_S;
  double* const Z = sm + L::O_Z;
  double* const IB = sm + L::O_IB;
  double* const T1 = sm + L::O_T1;
  double* const T2 = sm + L::O_T2;
  double* const RPE = sm + L::O_RPE;
  double* const SOL = sm + L::O_SOL;
  double* const RED = sm + L::O_RED;
  const double* const region = sm + L::O_REGION;
  const double tol2 = a.kkt_tol * a.kkt_tol, tau = 0.995;  // qp.cc:192

  DM_T_INIT();
  if (tid < 8) sm[L::O_ONES + tid] = 1.0;
  for (int i = tid; i < 8 * L::PLD; i += THREADS) sm[L::O_EYE + i] = (i / L::PLD == i % L::PLD) ? 1.0 : 0.0;
  for (long long prob = blockIdx.x; prob < a.batch; prob += gridDim.x) {
    const double* __restrict__ Gp = a.G + prob * (N * N);
    const double* __restrict__ Ap = K > 0 ? a.A + prob * (N * K) : nullptr;
    double* __restrict__ vp = a.vars + prob * STRIDE;
    __syncthreads();  // previous problem fully retired
    dm_stage_issue<N, K, W>(sm + L::O_REGION, Gp, Ap);
    for (int i = tid; i < N; i += THREADS) { sm[L::O_C + i] = a.c[prob * N + i]; X[i] = 0.0; }
    for (int i = tid; i < K; i += THREADS) { sm[L::O_B + i] = a.b[prob * K + i]; Y[i] = 0.0; }
    for (int i = tid; i < M; i += THREADS) IB[i] = a.ib[prob * M + i];
    double acc[MAXS][2];
    int status = -1, iters = 0;
    // ---- ComputeInitialGuess (qp.cc:439-482). SOLVE_EQUALITY_CONSTRAINED shares the factor/solve call site of the
    //      Newton loop below (guess_pending): the KKT system without Sigma, rhs = -[c; b_eq]. ----
    bool guess_pending = a.guess == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED;
    if (a.guess == MO_GUESS_USER_PROVIDED) {
      for (int i = tid; i < N; i += THREADS) X[i] = vp[i];
      for (int i = tid; i < M; i += THREADS) { S[i] = vp[N + i]; Z[i] = vp[N + M + K + i]; }
      for (int i = tid; i < K; i += THREADS) Y[i] = vp[N + M + i];
    } else if (!guess_pending) {
      __syncthreads();
      if (tid < N) dm_clamp_init(X, S, Z, IB, tid);
    }
    double mu = a.initial_mu, alpha_p = 1.0, alpha_d = 1.0;
    double kkt[4] = {0.0, 0.0, 0.0, 0.0};
    bool first = true;
    // ---- Solve loop (qp.cc:110-149); each trip = residuals of the current state, bookkeeping, one Newton step ----
    while (status < 0) {
      DM_T(10);
      cp_async_wait_all();
      __syncthreads();
      DM_T(11);
      double rhs = tid < N ? -sm[L::O_C + tid] : (tid < P ? -sm[L::O_B + tid - N] : 0.0);  // guess: -[c; b_eq]
      if (!guess_pending) {
      // (1) EvaluateKKTConditions (qp.cc:391-420): symmetric matvec from the staged lower trapezoid
      for (int task = tid; task < 2 * N + K; task += THREADS) {
        if (task < N) {  // row pass: c_i + sum_{j <= i} G(i,j) x_j
          const int i = task;
          double s0 = sm[L::O_C + i], s1 = 0.0;
          for (int jb = 0; 8 * jb <= i; ++jb) {
            const int len = N - 8 * jb + 2;
            const double* col = region + 8 * (jb * (N + 2) - 4 * jb * (jb - 1)) + i - 8 * jb;
            const int cnt = i - 8 * jb + 1 < 8 ? i - 8 * jb + 1 : 8;
#pragma unroll
            for (int jj = 0; jj < 8; jj += 2) {
              if (jj < cnt) s0 = fma(col[jj * len], X[8 * jb + jj], s0);
              if (jj + 1 < cnt) s1 = fma(col[(jj + 1) * len], X[8 * jb + jj + 1], s1);
            }
          }
          T1[i] = s0 + s1;
        } else if (task < 2 * N) {  // column pass: sum_{i > j} G(i,j) x_i  -  (A' y)_j
          const int j = task - N;
          const double* col = region + L::gcol_off(j) - 8 * (j / 8);
          double s0 = 0.0, s1 = 0.0;
          int i = j + 1;
          if ((i & 1) && i < N) { s0 = fma(col[i], X[i], s0); ++i; }
          for (; i + 1 < N; i += 2) { s0 = fma(col[i], X[i], s0); s1 = fma(col[i + 1], X[i + 1], s1); }
          if (i < N) s0 = fma(col[i], X[i], s0);
          if (K > 0) {
            const double* acol = region + L::GSIZE + j * L::ALD;
            double t0 = 0.0, t1 = 0.0;
#pragma unroll
            for (int r2 = 0; r2 < K; r2 += 2) { t0 = fma(acol[r2], Y[r2], t0); t1 = fma(acol[r2 + 1], Y[r2 + 1], t1); }
            s0 -= t0; s1 -= t1;
          }
          T2[j] = s0 + s1;
        } else {  // r_pe = A x + b
          const int r2 = task - 2 * N;
          const double* arow = region + L::GSIZE + r2;
          double s0 = sm[L::O_B + r2], s1 = 0.0;
#pragma unroll 8
          for (int j = 0; j < N; j += 2) { s0 = fma(arow[j * L::ALD], X[j], s0); s1 = fma(arow[(j + 1) * L::ALD], X[j + 1], s1); }
          RPE[r2] = s0 + s1;
        }
      }
      __syncthreads();
      // (2) per-variable residuals, Sigma, r_aug pieces; norms (ComputeErrors, qp.cc:423-437)
      double red[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // |r_d|^2, |r_pe|^2, |r_pi|^2, sum (sz)^2, sum sz, #(s <= 0)
      if (tid < N) {
        const int v = tid;
        const double base = T1[v] + T2[v];
        const double s0 = S[2 * v], s1 = S[2 * v + 1], z0 = Z[2 * v], z1 = Z[2 * v + 1], xv = X[v];
        const double p0 = (xv + IB[2 * v]) - s0, p1 = (IB[2 * v + 1] - xv) - s1;
        const double i0 = fast_rcp(s0), i1 = fast_rcp(s1);
        const double w0 = z0 * i0, w1 = z1 * i1;
        sm[L::O_SIG + v] = w0 + w1;
        const double rdv = (base - z0) + z1;
        const double c0 = s0 * z0, c1 = s1 * z1;
        red[0] = rdv * rdv;
        red[2] = fma(p0, p0, p1 * p1);
        red[3] = fma(c0, c0, c1 * c1);
        red[4] = c0 + c1;
        red[5] = ((s0 > 0.0) ? 0.0 : 1.0) + ((s1 > 0.0) ? 0.0 : 1.0);
        T1[v] = base + fma(w0, p0, -(w1 * p1));  // r_aug without the mu term (qp.cc:337-342)
        T2[v] = i0 - i1;
      } else if (tid < P) {
        const double pe = RPE[tid - N];
        red[1] = pe * pe;
      }
      DM_T(12);
      // The KKT assembly rides between the two barriers of the reduction: Sigma is visible after the first, and the
      // second guarantees every warp is done reading the staged G, A before the owner of tile row 0 writes L_00 over
      // them in the factorization prologue (a problem that terminates below wastes one assembly).
      if (tid == 0) *reinterpret_cast<int*>(sm + L::O_RED + 8 * W) = 0;  // zero-pivot flag of the coming factorization
      block_reduce<W, 6, false>(red, RED, [&] { dm_assemble<N, K, W, MAXS>(acc, sm, true, tid); });
      DM_T(13);
      // (3) bookkeeping of the Solve loop: errors of the step just taken, termination, barrier update
      const double mu_c = red[4] / (double)M;  // ComputeMu, qp.cc:509-516
      const double corrected = red[3] - 2 * (red[4] * mu) + (mu * mu) * M;
      const double comp2 = corrected > 0.0 ? corrected : 0.0;
      if (first) {
        first = false;
        mu = a.init_mu_with_comp ? mu_c : a.initial_mu;
      } else {
        const double max2 = fmax(fmax(red[0], comp2), fmax(red[1], red[2]));
        kkt[0] = sqrt(red[0]); kkt[1] = sqrt(comp2); kkt[2] = sqrt(red[1]); kkt[3] = sqrt(red[2]);
        if (a.iter_log && tid == 0) {
          double* row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
          row[4] = kkt[0]; row[5] = kkt[1]; row[6] = kkt[2]; row[7] = kkt[3];
          row[8] = mu; row[9] = alpha_p; row[10] = alpha_d; row[11] = qnan(); row[12] = qnan(); row[13] = qnan();
        }
        ++iters;
        if (max2 < tol2 && mu_c < a.comp_tol) status = MO_QP_SATISFIED_KKT_TOL;  // qp.cc:132-137
        else if (iters >= a.max_iterations) status = MO_QP_MAX_ITERATIONS;
        else if (max2 <= mu * mu || !a.decrease_mu_only_on_small_error)          // qp.cc:140-146
          mu = (a.strategy == MO_BARRIER_FIXED_DECREASE) ? mu * a.sigma : a.sigma * mu_c;
      }
      if (status < 0 && red[5] != 0.0) status = MO_QP_NONPOSITIVE_SLACK;  // F_ASSERT(all s > 0), qp.cc:285
      if (status >= 0) break;
      if (a.iter_log && tid == 0) {  // kkt_initial of the coming iteration (ComputeErrors with the new mu, qp.cc:118)
        double* row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
        const double cr = red[3] - 2 * (red[4] * mu) + (mu * mu) * M;
        row[0] = sqrt(red[0]); row[1] = sqrt(cr > 0.0 ? cr : 0.0); row[2] = sqrt(red[1]); row[3] = sqrt(red[2]);
      }
      // (4) ComputeLDLT + SolveForUpdate: rhs = -[r_aug; r_pe]  ->  SOL = [dx; -dy]
        rhs = tid < N ? -fma(-mu, T2[tid], T1[tid]) : (tid < P ? -RPE[tid - N] : 0.0);
      }
      DM_T(14);
      if (guess_pending) {  // KKT system without Sigma (the Newton trips assemble inside the reduction above)
        if (tid == 0) *reinterpret_cast<int*>(sm + L::O_RED + 8 * W) = 0;
        dm_assemble<N, K, W, MAXS>(acc, sm, false, tid);
        __syncthreads();
      }
      DM_T(15);
      const bool ok = dm_factor_solve<N, K, W, MAXS>(acc, rhs, sm, tid);
      if (!ok) { status = MO_QP_FAILED_FACTORIZATION; break; }
      dm_stage_issue<N, K, W>(sm + L::O_REGION, Gp, Ap);  // next trip's G, A: overlaps the update below
      if (guess_pending) {  // x <- dx, y <- dy of the equality constrained solve (qp.cc:452-459), then clamp / s / z
        guess_pending = false;
        if (tid < N) { X[tid] = SOL[tid]; dm_clamp_init(X, S, Z, IB, tid); }
        else if (tid < P) Y[tid - N] = -SOL[tid];
        continue;
      }
      // (5) ds, dz (qp.cc:359-363), step lengths (:485-507), state update (:196-199)
      double al[2] = {1.0, 1.0};
      double d0 = 0.0, d1 = 0.0, e0 = 0.0, e1 = 0.0;
      if (tid < N) {
        const int v = tid;
        const double s0 = S[2 * v], s1 = S[2 * v + 1], z0 = Z[2 * v], z1 = Z[2 * v + 1], xv = X[v], dxv = SOL[v];
        const double p0 = (xv + IB[2 * v]) - s0, p1 = (IB[2 * v + 1] - xv) - s1;
        const double i0 = fast_rcp(s0), i1 = fast_rcp(s1);
        d0 = dxv + p0; d1 = p1 - dxv;
        e0 = fma(mu, i0, -fma(z0 * i0, d0, z0));
        e1 = fma(mu, i1, -fma(z1 * i1, d1, z1));
        if (s0 + d0 <= 0.0 && fabs(d0) > 0) al[0] = fmin(al[0], -tau * s0 / d0);
        if (s1 + d1 <= 0.0 && fabs(d1) > 0) al[0] = fmin(al[0], -tau * s1 / d1);
        if (z0 + e0 <= 0.0 && fabs(e0) > 0) al[1] = fmin(al[1], -tau * z0 / e0);
        if (z1 + e1 <= 0.0 && fabs(e1) > 0) al[1] = fmin(al[1], -tau * z1 / e1);
      }
      block_reduce<W, 2, true>(al, RED);
      alpha_p = al[0]; alpha_d = al[1];
      if (tid < N) {
        const int v = tid;
        X[v] = fma(SOL[v], alpha_p, X[v]);
        S[2 * v] = fma(d0, alpha_p, S[2 * v]); S[2 * v + 1] = fma(d1, alpha_p, S[2 * v + 1]);
        Z[2 * v] = fma(e0, alpha_d, Z[2 * v]); Z[2 * v + 1] = fma(e1, alpha_d, Z[2 * v + 1]);
      } else if (tid < P) {
        Y[tid - N] = fma(-SOL[tid], alpha_d, Y[tid - N]);
      }
    }
    // ---- results ----
    cp_async_wait_all();
    __syncthreads();
    for (int i = tid; i < N; i += THREADS) vp[i] = X[i];
    for (int i = tid; i < M; i += THREADS) { vp[N + i] = S[i]; vp[N + M + K + i] = Z[i]; }
    for (int i = tid; i < K; i += THREADS) vp[N + M + i] = Y[i];
    if (tid == 0) {
      a.status[prob] = status;
      a.iters[prob] = iters;
      if (a.kkt_final) {
        double* kf = a.kkt_final + prob * 4;
        kf[0] = kkt[0]; kf[1] = kkt[1]; kf[2] = kkt[2]; kf[3] = kkt[3];
      }
      if (a.lagrange) {  // ComputeLagrangeMultiplierSummary, qp.cc:539-546
        double mn = qnan(), li = qnan();
        if (K > 0) {
          mn = Y[0]; li = fabs(Y[0]);
          for (int r2 = 1; r2 < K; ++r2) { mn = fmin(mn, Y[r2]); li = fmax(li, fabs(Y[r2])); }
        }
        a.lagrange[prob * 2] = mn;
        a.lagrange[prob * 2 + 1] = li;
      }
    }
  }
  DM_T_DUMP();
}

}  // namespace mo
