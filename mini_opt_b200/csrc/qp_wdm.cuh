// WARP-per-problem interior point QP solve for medium dense box-constrained QPs (n, k multiples of 8, n + k <= 80):
// the reduced KKT matrix lives ONLY in shared memory as 8x8 tiles, the factorization is a LEFT-looking unpivoted LDL'
// whose O(p^3) work is mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4), and every vector of the iteration lives in registers.
//
// Replaces, per problem: QPInteriorPointSolver::Setup + Solve (source/qp.cc:20-73, 100-151): ComputeInitialGuess
// (:439-482), the Iterate loop (:153-201), EvaluateKKTConditions (:391-420), ComputeLDLT (:275-316) + SolveForUpdate
// (:318-364) as ONE factorization and ONE two-sided solve, ComputeAlpha (:485-507), ComputeMu (:509-516),
// ComputeErrors (:423-437).
//
// Why a second DMMA kernel next to qp_dm_kernel (CTA per problem, matrix in accumulator registers AND L in shared
// memory): one factorization is a chain of p dependent pivots (~150-200 cycles each), so a CTA-per-problem kernel is
// bound by that latency times the number of co-resident problems, and two on-chip copies of the matrix cap residency
// at 4 problems per SM (n = 64). Here a problem costs one warp and ONE copy of the matrix (28 KB at n = 64, 10 KB at
// n = 32): 7 resp. 12 independent problems per SM (shared memory resp. a spill-free register budget), no CTA barriers,
// and the pivot chains of different problems overlap.
//
// Layout: tile (I, J), I >= J, 64 doubles, element [r][c] at wd_el(r, c) = (c >> 2) * 32 + r' * 4 + (c & 3) with r' = r
// (first half) or r ^ 2 (second half): lane l = 4 r + q of a warp reads [r][4h + q] at l resp. 32 + (l ^ 8) -- which is at once the DMMA A fragment of the tile and the B fragment of its
// transpose, 256 contiguous bytes, bank-conflict free. For n <= 32 tile row NT (one past the matrix) carries the
// right-hand side in its row 0, so the forward substitution and the D^-1 scaling fall out of the panel solve; at n = 64
// the rhs stays out of the tile array (WdLayout::RHS_ROW, wd_forward_block).
#pragma once
#include <cstdio>
#include "qp_dmma.cuh"

namespace mo {

template <int N, int K>
struct WdLayout {
  // RHS_ROW: the right-hand side rides through the factorization as tile row NT (forward substitution for free).
  // At n = 64 that row costs 5 KB and the 7th resident problem per SM; there the forward substitution is done
  // block by block next to the diagonal block instead (wd_forward_block).
  static constexpr bool RHS_ROW = N < 64;
  static constexpr int P = N + K, NT = P / 8, NTN = N / 8, NTR = NT + (RHS_ROW ? 1 : 0), M = 2 * N;
  static constexpr int VPL = (N + 31) / 32;  // variables per lane
  // stage the next trip's G, A column by column inside the back substitution (measured: n = 64 19.85 -> 19.53 ms;
  // n = 32 / 16 slower, their loop is issue-bound) or in one piece after it
  static constexpr bool STAGE_IN_BACKSUB = N >= 64;
  static_assert(N % 8 == 0 && K % 8 == 0 && N > 0 && K > 0 && K <= 32, "tile shapes");
  __host__ __device__ static constexpr int colbase(int J) { return J * NTR - J * (J - 1) / 2; }
  __host__ __device__ static constexpr int tile(int I, int J) { return 64 * (colbase(J) + (I - J)); }
  static constexpr int NTILES = colbase(NT);
  static constexpr int O_D = 64 * NTILES;   // pivots d[P]
  static constexpr int O_INVD = O_D + P;    // 1 / d
  static constexpr int O_XV = O_INVD + P;   // [x; -y] for the matvec, then the solution [dx; -dy]
  static constexpr int O_R = O_XV + P;      // K [x; -y]; !RHS_ROW: then the right-hand side
  static constexpr int O_W = O_R + P;       // !RHS_ROW: w = D^-1 L^-1 rhs [P]
  static constexpr int O_RC = O_W;          // two-warp variant: the helper warp's column sums of the matvec [N] (aliases
                                            // w: written in the matvec phase, w lives from factorization to back substitution)
  static constexpr int O_CTL = O_W + P;     // two-warp variant: status broadcast
  static constexpr int TOTAL = O_CTL + 2;
  static constexpr size_t SMEM_BYTES = sizeof(double) * TOTAL;
};

#ifdef MO_WD_TRACE
// Debug build only: lane 0 of CTA 0 accumulates cycles per phase in shared memory and prints them at kernel end.
__shared__ long long s_wd_tr[16];
__shared__ long long s_wd_t0;
#define WD_T(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long t_ = clock64(); s_wd_tr[i] += t_ - s_wd_t0; s_wd_t0 = t_; } } while (0)
#else
#define WD_T(i) do {} while (0)
#endif

// Offset of element [r][c] inside a tile: two 8x4 halves (columns 0-3, 4-7) of 32 doubles, row-major; the rows of the
// second half are XOR-swizzled by 2 so that the 16-byte accumulator-fragment accesses of a quarter warp (rows 2a, 2a+1
// of BOTH halves) fall into 32 distinct banks (unswizzled: 2-way conflicts on every accumulator load / store, 13 % of
// all shared-memory wavefronts with the pipe 62 % busy, profiles/r01_wd64_ncu_full_summary.txt).
__host__ __device__ __forceinline__ constexpr int wd_el(int r, int c) {
  return (c >> 2) * 32 + (((c >> 2) ? (r ^ 2) : r) << 2) + (c & 3);
}

__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
  const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gsrc) : "memory");
}

__device__ __forceinline__ double wd_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double wd_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// sum over the 8 lanes that share q = lane & 3 (lanes differ in r = lane >> 2)
__device__ __forceinline__ double wd_sum_r(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  return v;
}

// LOCK-STEP CTAs (LOCK > 1): LOCK warps per CTA, each with its own problem in its own shared-memory region, walk the block
// columns of their factorizations TOGETHER -- a CTA barrier before and after every diagonal block -- so that an SM is either
// in a DMMA phase (panel solve of column J + gather of column J+1, all warps) or in a pivot-chain phase (diagonal block J,
// all warps). Why: DMMA.8x8x4 holds the FP64 pipe of a sub-partition for 16 cycles; with the phases of the resident warps
// interleaved at random every dependent DFMA of a pivot chain queues behind other warps' DMMAs (16-25 cycles per dependent
// operation instead of 8, profiles/r02_wd_variants.txt, last lines: tools/ubench/lanes_bench.cu), and the chain is 37 % of the kernel
// (profiles/r02_wd64_phase_trace.txt). The barriers order nothing -- no data is shared between the warps.
// MEASURED (profiles/r02_wd_lockstep_ab.txt): the hypothesis is wrong for this kernel. With every warp of the SM in its
// pivot-chain phase at once the diagonal blocks take exactly as long (1.60M cycles of a 4.4M-cycle launch either way): the
// block costs ~430 issued instructions at ~3.7 cycles each, not queueing behind DMMAs; the barriers only add waiting
// (n=64 19.2 vs 18.7 ms, n=32 23.5 vs 18.2, n=16 25.9 vs 18.5). Off by default, kept behind mo_debug_set_wd_lock.
template <int LOCK>
__device__ __forceinline__ void wd_lock_bar() {
  if constexpr (LOCK > 1) __syncthreads();
}
template <int N, int K>
__host__ __device__ constexpr int wd_lock_barriers() { return 2 * WdLayout<N, K>::NT; }  // per factorization, after the opening __syncthreads_or

// cp.async staging of G (lower block triangle) and A_eq into the tiles; every request is 8 rows x 4 columns:
// four 64-byte column segments in global memory, 256 contiguous bytes in shared memory.
template <int N, int K, int NW>
__device__ __forceinline__ void wd_stage(double* __restrict__ T, const double* __restrict__ Gp,
                                         const double* __restrict__ Ap, int lane, int warp) {
  using L = WdLayout<N, K>;
  const int r = lane >> 2, q = lane & 3;
  const double* __restrict__ g = Gp + q * N + r;
  const double* __restrict__ a = Ap + q * K + r;
  int cnt = 0;  // the warps of a problem take alternate tiles
#pragma unroll
  for (int J = 0; J < L::NTN; ++J) {
#pragma unroll
    for (int I = J; I < L::NTN; ++I, ++cnt) {
      if (NW == 1 || (cnt % NW) == warp) {
        cp_async8(T + L::tile(I, J) + lane, g + (8 * J) * N + 8 * I);
        cp_async8(T + L::tile(I, J) + 32 + (lane ^ 8), g + (8 * J + 4) * N + 8 * I);
      }
    }
#pragma unroll
    for (int I = L::NTN; I < L::NT; ++I, ++cnt) {
      if (NW == 1 || (cnt % NW) == warp) {
        cp_async8(T + L::tile(I, J) + lane, a + (8 * J) * K + 8 * (I - L::NTN));
        cp_async8(T + L::tile(I, J) + 32 + (lane ^ 8), a + (8 * J + 4) * K + 8 * (I - L::NTN));
      }
    }
  }
  cp_async_commit();
}

// LDL' of the 8x8 diagonal block stored in `tile` (lower triangle significant), computed REDUNDANTLY by every lane
// from a register copy of the block: no shuffles and no selects on the pivot chain, which is the critical path of the
// whole factorization (a lane-per-row version with shuffles measured 1460 cycles per block, its inverse 640). Pivots in
// PAIRS (1/d1 = a00 / det, the multiplication folded into the last step of the reciprocal) so that the two reciprocal
// chains of a pair overlap; the entries that feed the NEXT pair are updated with a single operation after 1/d1
// arrives. Writes d, 1/d and OVERWRITES the tile with S = inv(L_JJ) (unit lower; the panel solve and the back
// substitution only ever need the inverse): lane c < 8 computes column c of S from the register copy of L.
// `chk` turns NaN on a zero or non-finite pivot.
__device__ __noinline__ void wd_diag_block(double* __restrict__ tile, double* __restrict__ d, double* __restrict__ invd,
                                           double& chk_io) {
  const int lane = threadIdx.x & 31, lr = lane & 7;
  double A[8][8];  // lower triangle; after the factorization the strict lower triangle holds L
#pragma unroll
  for (int i = 0; i < 8; ++i) {
#pragma unroll
    for (int j = 0; j <= i; j += 2) {
      const double2 v = *reinterpret_cast<const double2*>(tile + wd_el(i, j));
      A[i][j] = v.x;
      if (j + 1 <= i) A[i][j + 1] = v.y;
    }
  }
  double chk = chk_io, dd[8], di[8];
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    const double a00 = A[c][c], a10 = A[c + 1][c], a11 = A[c + 1][c + 1];
    const double det = fma(a00, a11, -(a10 * a10));
    const double inv0 = rcp_chain(a00);
    double xd;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(xd) : "d"(det));
    const double ed = fma(-det, xd, 1.0), yd = a00 * xd;
    const double inv1 = fma(yd, fma(ed, ed, ed), yd);  // 1 / d1, d1 = det / a00
    dd[c] = a00; dd[c + 1] = fma(-(a10 * inv0), a10, a11);
    di[c] = inv0; di[c + 1] = inv1;
    chk = fma(inv0, 0.0, chk);
    chk = fma(inv1, 0.0, chk);
    A[c + 1][c] = a10 * inv0;
    double u0[8], u1[8], l0[8];
#pragma unroll
    for (int i = c + 2; i < 8; ++i) {
      u0[i] = A[i][c];
      l0[i] = u0[i] * inv0;
      u1[i] = fma(-l0[i], a10, A[i][c + 1]);
    }
#pragma unroll
    for (int i = c + 2; i < 8; ++i) {
#pragma unroll
      for (int j = c + 2; j <= i; ++j)  // A(i,j) -= l0_i u0_j + (u1_i u1_j) / d1
        A[i][j] = fma(-(u1[i] * u1[j]), inv1, fma(-l0[i], u0[j], A[i][j]));
      A[i][c] = l0[i];
      A[i][c + 1] = u1[i] * inv1;
    }
  }
  chk_io = chk;
  // column lr of inv(L): x_i = e_i - sum_{k < i} L(i,k) x_k (x_i = 0 above the diagonal, 1 on it)
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    double v = i == lr ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < i; ++k) v = fma(-A[i][k], x[k], v);
    x[i] = i >= lr ? v : 0.0;
  }
  double myd = dd[0], myinv = di[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) { myd = lr == i ? dd[i] : myd; myinv = lr == i ? di[i] : myinv; }
  __syncwarp();  // every lane has read the block
  if (lane < 8) {
    d[lr] = myd; invd[lr] = myinv;
#pragma unroll
    for (int i = 0; i < 8; ++i) tile[wd_el(i, lr)] = x[i];  // S[i][lr]
  }
  __syncwarp();
}

// Forward substitution of block J when the right-hand side is not a tile row (WdLayout::RHS_ROW == false):
// w_J = D_J^-1 S_J (b_J - sum_{kk < J} L(J,kk) D_kk w_kk), b in sm[O_R], w in sm[O_W]. Row sums over q (2 shuffle
// levels), then S_J read transposed so that t[r] is already where it is needed (3 levels).
template <int N, int K, int J>
__device__ __forceinline__ void wd_forward_block(double* __restrict__ sm, int lane) {
  using L = WdLayout<N, K>;
  const int r = lane >> 2, q = lane & 3;
  const double* __restrict__ T = sm;
  const double* __restrict__ D = sm + L::O_D;
  double* __restrict__ W = sm + L::O_W;
  double sa = 0.0, sb = 0.0, sc = 0.0, sd = 0.0;
#pragma unroll
  for (int kk = 0; kk < J; ++kk) {
    const double* __restrict__ t = T + L::tile(J, kk);
    const double z0 = W[8 * kk + q] * D[8 * kk + q], z1 = W[8 * kk + 4 + q] * D[8 * kk + 4 + q];
    if (kk & 1) { sc = fma(t[lane], z0, sc); sd = fma(t[32 + (lane ^ 8)], z1, sd); }
    else { sa = fma(t[lane], z0, sa); sb = fma(t[32 + (lane ^ 8)], z1, sb); }
  }
  double s = (sa + sb) + (sc + sd);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  const double tr = sm[L::O_R + 8 * J + r] - s;  // t[r] (replicated over q)
  const int toff = wd_el(q, r);  // element [q][r]; [4 + q][r] is 16 further
  const double* __restrict__ S = T + L::tile(J, J);
  double z0 = S[toff] * tr, z1 = S[toff + 16] * tr;  // S[4h + q][r] t[r]
  z0 = wd_sum_r(z0); z1 = wd_sum_r(z1);
  if (r == 0) {
    W[8 * J + q] = z0 * sm[L::O_INVD + 8 * J + q];
    W[8 * J + 4 + q] = z1 * sm[L::O_INVD + 8 * J + 4 + q];
  }
  __syncwarp();
}

// Block column J of the left-looking factorization: gather the updates of all earlier panels into the accumulator
// fragments of tiles (J..NT, J), factor the diagonal block, solve the panel (and the right-hand-side row) with DMMA
// against T = inv(L_JJ)' D_J^-1.
template <int N, int K, int J, int LOCK = 1>
__device__ __forceinline__ void wd_column(double* __restrict__ sm, int lane, double& chk) {
  using L = WdLayout<N, K>;
  constexpr int CNT = L::NTR - J;
  double* __restrict__ T = sm;
  const int r = lane >> 2, q = lane & 3;
  const int coff = wd_el(r, 2 * q);  // accumulator fragment: [r][2q], [r][2q+1]
  double acc[CNT][2];
#pragma unroll
  for (int i = 0; i < CNT; ++i) {
    const double2 v = *reinterpret_cast<const double2*>(T + L::tile(J + i, J) + coff);
    acc[i][0] = v.x; acc[i][1] = v.y;
  }
  if constexpr (J > 0) {
    // fully unrolled: tile addresses are immediates and the scheduler hoists the fragment loads of panel kk+1 above
    // the DMMAs of panel kk (a runtime loop spent more cycles on addressing and load latency than on DMMA)
#pragma unroll
    for (int kk = 0; kk < J; ++kk) {
      const double* __restrict__ tb = T + L::tile(J, kk);
      const double b0 = tb[lane] * -sm[L::O_D + 8 * kk + q];
      const double b1 = tb[32 + (lane ^ 8)] * -sm[L::O_D + 8 * kk + 4 + q];
      double a0[CNT], a1[CNT];
#pragma unroll
      for (int i = 0; i < CNT; ++i) { a0[i] = tb[64 * i + lane]; a1[i] = tb[64 * i + 32 + (lane ^ 8)]; }
#pragma unroll
      for (int i = 0; i < CNT; ++i) dmma884(acc[i][0], acc[i][1], a0[i], b0);
#pragma unroll
      for (int i = 0; i < CNT; ++i) dmma884(acc[i][0], acc[i][1], a1[i], b1);
    }
  }
#pragma unroll
  for (int i = 0; i < CNT; ++i)
    *reinterpret_cast<double2*>(T + L::tile(J + i, J) + coff) = make_double2(acc[i][0], acc[i][1]);
  __syncwarp();
  wd_lock_bar<LOCK>();  // DMMA phase of the CTA ends
  WD_T(2);
  wd_diag_block(T + L::tile(J, J), sm + L::O_D + 8 * J, sm + L::O_INVD + 8 * J, chk);
  if constexpr (!L::RHS_ROW) wd_forward_block<N, K, J>(sm, lane);
  wd_lock_bar<LOCK>();  // pivot-chain phase ends
  WD_T(3);
  const double idr = sm[L::O_INVD + 8 * J + r];
  const double bt0 = T[L::tile(J, J) + lane] * idr, bt1 = T[L::tile(J, J) + 32 + (lane ^ 8)] * idr;
  double a0[CNT], a1[CNT];
#pragma unroll
  for (int i = 1; i < CNT; ++i) { a0[i] = T[L::tile(J + i, J) + lane]; a1[i] = T[L::tile(J + i, J) + 32 + (lane ^ 8)]; }
  __syncwarp();
#pragma unroll
  for (int i = 1; i < CNT; ++i) {
    double p0 = 0.0, p1 = 0.0, q0 = 0.0, q1 = 0.0;
    dmma884(p0, p1, a0[i], bt0);
    dmma884(q0, q1, a1[i], bt1);
    *reinterpret_cast<double2*>(T + L::tile(J + i, J) + coff) = make_double2(p0 + q0, p1 + q1);
  }
  __syncwarp();
  WD_T(4);
}

// Two-warp variant of the block column step (look-ahead): warp 0 factors diagonal block J and solves panel J while
// warp 1 gathers block column J+1 over the panels kk < J (independent of block J); after a barrier warp 1 applies the
// kk = J update and spills column J+1. The pivot chain of block J and the DMMA gather of column J+1 overlap in
// hardware instead of queueing in one in-order instruction stream.
template <int N, int K, int J>
__device__ __forceinline__ void wd2_column(double* __restrict__ sm, int lane, int warp, double& chk) {
  using L = WdLayout<N, K>;
  constexpr int CP = L::NTR - J - 1;      // tiles below the diagonal block: I = J+1 .. NT (= the tiles of column J+1)
  constexpr bool HASN = J + 1 < L::NT;    // a block column J+1 exists
  double* __restrict__ T = sm;
  const int r = lane >> 2, q = lane & 3;
  const int coff = wd_el(r, 2 * q);
  constexpr int CPA = CP > 0 ? CP : 1;
  double acc[CPA][2];
  if (warp == 0) {
    wd_diag_block(T + L::tile(J, J), sm + L::O_D + 8 * J, sm + L::O_INVD + 8 * J, chk);
    if constexpr (!L::RHS_ROW) wd_forward_block<N, K, J>(sm, lane);
    const double idr = sm[L::O_INVD + 8 * J + r];
    const double bt0 = T[L::tile(J, J) + lane] * idr, bt1 = T[L::tile(J, J) + 32 + (lane ^ 8)] * idr;
    double f0[CPA], f1[CPA];
#pragma unroll
    for (int i = 0; i < CP; ++i) { f0[i] = T[L::tile(J + 1 + i, J) + lane]; f1[i] = T[L::tile(J + 1 + i, J) + 32 + (lane ^ 8)]; }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < CP; ++i) {
      double p0 = 0.0, p1 = 0.0, q0 = 0.0, q1 = 0.0;
      dmma884(p0, p1, f0[i], bt0);
      dmma884(q0, q1, f1[i], bt1);
      *reinterpret_cast<double2*>(T + L::tile(J + 1 + i, J) + coff) = make_double2(p0 + q0, p1 + q1);
    }
  } else if constexpr (HASN) {
#pragma unroll
    for (int i = 0; i < CP; ++i) {
      const double2 v = *reinterpret_cast<const double2*>(T + L::tile(J + 1 + i, J + 1) + coff);
      acc[i][0] = v.x; acc[i][1] = v.y;
    }
#pragma unroll
    for (int kk = 0; kk < J; ++kk) {
      const double* __restrict__ tb = T + L::tile(J + 1, kk);
      const double b0 = tb[lane] * -sm[L::O_D + 8 * kk + q];
      const double b1 = tb[32 + (lane ^ 8)] * -sm[L::O_D + 8 * kk + 4 + q];
      double a0[CPA], a1[CPA];
#pragma unroll
      for (int i = 0; i < CP; ++i) { a0[i] = tb[64 * i + lane]; a1[i] = tb[64 * i + 32 + (lane ^ 8)]; }
#pragma unroll
      for (int i = 0; i < CP; ++i) dmma884(acc[i][0], acc[i][1], a0[i], b0);
#pragma unroll
      for (int i = 0; i < CP; ++i) dmma884(acc[i][0], acc[i][1], a1[i], b1);
    }
  }
  WD_T(3);
  __syncthreads();
  WD_T(2);
  if constexpr (HASN) {
    if (warp == 1) {
      const double* __restrict__ tb = T + L::tile(J + 1, J);
      const double b0 = tb[lane] * -sm[L::O_D + 8 * J + q];
      const double b1 = tb[32 + (lane ^ 8)] * -sm[L::O_D + 8 * J + 4 + q];
      double a0[CPA], a1[CPA];
#pragma unroll
      for (int i = 0; i < CP; ++i) { a0[i] = tb[64 * i + lane]; a1[i] = tb[64 * i + 32 + (lane ^ 8)]; }
#pragma unroll
      for (int i = 0; i < CP; ++i) {
        double e0 = 0.0, e1 = 0.0;  // second k-half in its own accumulator: two independent DMMAs per tile
        dmma884(acc[i][0], acc[i][1], a0[i], b0);
        dmma884(e0, e1, a1[i], b1);
        *reinterpret_cast<double2*>(T + L::tile(J + 1 + i, J + 1) + coff) = make_double2(acc[i][0] + e0, acc[i][1] + e1);
      }
    }
    __syncthreads();
    WD_T(4);
  }
}


// ---------------------------------------------------------------------------------------------------------
// LOOK-AHEAD schedule of the block column step (LA = true), the default at n = 64.
//
// Measured motivation (profiles/r02_wd64_phase_trace.txt, 7 warps per SM): with the plain left-looking order a column
// is gather (1.2k cycles) -> diagonal block + forward substitution (2.4k, a pure latency chain: 4 pivot pairs of ~9
// dependent FP64 operations each, every one of them queueing behind the 16-cycle DMMAs of the other warps of the SM
// sub-partition) -> panel solve (0.5k), strictly one after the other in the instruction stream of the one warp that
// owns the problem: 37 % of the kernel is the diagonal chain, during which the warp issues one instruction every ~25 cycles.
// Here the chain runs UNDER the DMMA work of the same column:
//   * invariant on entry: diagonal tile (J, J) is completely updated (all panels kk < J applied);
//   * one basic block holds (a) the register copy + pivot chain of block J and (b) the gather of the off-diagonal
//     tiles (J+1.., J) over kk < J together with the look-ahead gather of the NEXT diagonal tile (J+1, J+1) over
//     kk < J -- (b) does not depend on (a), so the scheduler fills the chain's latency slots with DMMA / LDS;
//   * the gathered tiles stay in their accumulator registers until the panel solve: C-fragment -> A-fragment is four
//     quad shuffles per tile (wdl_c2a) instead of a store / barrier / load round trip through shared memory;
//   * after the panel solve of tile (J+1, J) the kk = J update goes straight into the look-ahead accumulator and the
//     finished diagonal tile (J+1, J+1) is stored: the next column starts with its chain at once;
//   * the forward substitution of the right-hand side (n = 64) is split: its dot products run in (b), its S_J product
//     after the panel solve, off the factorization's critical path.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void wdl_c2a(double c0, double c1, int lane, double& a0, double& a1) {
  // lane 4r + q holds [r][2q], [r][2q+1]; wants [r][q], [r][4 + q]: column q sits in lane 4r + (q >> 1), slot q & 1
  const int src = (lane & ~3) + ((lane & 3) >> 1);
  const double t0 = __shfl_sync(0xffffffffu, c0, src), t1 = __shfl_sync(0xffffffffu, c1, src);
  const double u0 = __shfl_sync(0xffffffffu, c0, src + 2), u1 = __shfl_sync(0xffffffffu, c1, src + 2);
  a0 = (lane & 1) ? t1 : t0;
  a1 = (lane & 1) ? u1 : u0;
}

template <int N, int K, int J>
__device__ __forceinline__ void wdl_column(double* __restrict__ sm, const int lane, double& chk_io) {
  using L = WdLayout<N, K>;
  constexpr int CNT1 = L::NTR - 1 - J;  // off-diagonal tiles of block column J (the rhs tile row included)
  constexpr bool HASD = J + 1 < L::NT;  // a next diagonal tile exists
  constexpr int CA = CNT1 > 0 ? CNT1 : 1;
  constexpr bool FWD = !L::RHS_ROW;     // block forward substitution of the rhs vector
  double* __restrict__ T = sm;
  const int r = lane >> 2, q = lane & 3, lr = lane & 7;
  const int coff = wd_el(r, 2 * q);
  const int hi = 32 + (lane ^ 8);
  // ---- (a1) register copy of the finished diagonal block ----
  double A[8][8];
  {
    const double* __restrict__ tile = T + L::tile(J, J);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int j = 0; j <= i; j += 2) {
        const double2 v = *reinterpret_cast<const double2*>(tile + wd_el(i, j));
        A[i][j] = v.x;
        if (j + 1 <= i) A[i][j + 1] = v.y;
      }
    }
  }
  // ---- (b) gather: off-diagonal tiles of this column and the next diagonal tile, panels kk < J ----
  double acc[CA][2], accD[2] = {0.0, 0.0};
#pragma unroll
  for (int i = 0; i < CNT1; ++i) {
    const double2 v = *reinterpret_cast<const double2*>(T + L::tile(J + 1 + i, J) + coff);
    acc[i][0] = v.x; acc[i][1] = v.y;
  }
  if constexpr (HASD) {
    const double2 v = *reinterpret_cast<const double2*>(T + L::tile(J + 1, J + 1) + coff);
    accD[0] = v.x; accD[1] = v.y;
  }
  double fsa = 0.0, fsb = 0.0, fsc = 0.0, fsd = 0.0;  // forward substitution: sum_kk L(J,kk) D_kk w_kk (row sums over q)
  if constexpr (J > 0) {
#pragma unroll
    for (int kk = 0; kk < J; ++kk) {
      const double* __restrict__ tb = T + L::tile(J, kk);
      const double dq0 = sm[L::O_D + 8 * kk + q], dq1 = sm[L::O_D + 8 * kk + 4 + q];
      const double l0 = tb[lane], l1 = tb[hi];
      const double b0 = l0 * -dq0, b1 = l1 * -dq1;
      if constexpr (FWD) {
        const double z0 = sm[L::O_W + 8 * kk + q] * dq0, z1 = sm[L::O_W + 8 * kk + 4 + q] * dq1;
        if (kk & 1) { fsc = fma(l0, z0, fsc); fsd = fma(l1, z1, fsd); }
        else { fsa = fma(l0, z0, fsa); fsb = fma(l1, z1, fsb); }
      }
      if constexpr (CNT1 > 0) {
        double a0[CA], a1[CA];
#pragma unroll
        for (int i = 0; i < CNT1; ++i) { a0[i] = tb[64 * (i + 1) + lane]; a1[i] = tb[64 * (i + 1) + hi]; }
        if constexpr (HASD) {  // tile (J+1, kk) is both operands of the diagonal update
          const double bd0 = a0[0] * -dq0, bd1 = a1[0] * -dq1;
          dmma884(accD[0], accD[1], a0[0], bd0);
#pragma unroll
          for (int i = 0; i < CNT1; ++i) dmma884(acc[i][0], acc[i][1], a0[i], b0);
          dmma884(accD[0], accD[1], a1[0], bd1);
#pragma unroll
          for (int i = 0; i < CNT1; ++i) dmma884(acc[i][0], acc[i][1], a1[i], b1);
        } else {
#pragma unroll
          for (int i = 0; i < CNT1; ++i) dmma884(acc[i][0], acc[i][1], a0[i], b0);
#pragma unroll
          for (int i = 0; i < CNT1; ++i) dmma884(acc[i][0], acc[i][1], a1[i], b1);
        }
      }
    }
  }
  // ---- (a2) pivot chain of the 8x8 block, redundantly in every lane (see wd_diag_block) ----
  double chk = chk_io, dd[8], di[8];
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    const double a00 = A[c][c], a10 = A[c + 1][c], a11 = A[c + 1][c + 1];
    const double det = fma(a00, a11, -(a10 * a10));
    const double inv0 = rcp_chain(a00);
    double xd;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(xd) : "d"(det));
    const double ed = fma(-det, xd, 1.0), yd = a00 * xd;
    const double inv1 = fma(yd, fma(ed, ed, ed), yd);  // 1 / d1, d1 = det / a00
    dd[c] = a00; dd[c + 1] = fma(-(a10 * inv0), a10, a11);
    di[c] = inv0; di[c + 1] = inv1;
    chk = fma(inv0, 0.0, chk);
    chk = fma(inv1, 0.0, chk);
    A[c + 1][c] = a10 * inv0;
    double u0[8], u1[8], l0[8];
#pragma unroll
    for (int i = c + 2; i < 8; ++i) {
      u0[i] = A[i][c];
      l0[i] = u0[i] * inv0;
      u1[i] = fma(-l0[i], a10, A[i][c + 1]);
    }
#pragma unroll
    for (int i = c + 2; i < 8; ++i) {
#pragma unroll
      for (int j = c + 2; j <= i; ++j) A[i][j] = fma(-(u1[i] * u1[j]), inv1, fma(-l0[i], u0[j], A[i][j]));
      A[i][c] = l0[i];
      A[i][c + 1] = u1[i] * inv1;
    }
  }
  chk_io = chk;
  double x[8];  // column lr of inv(L_JJ)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    double v = i == lr ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < i; ++k) v = fma(-A[i][k], x[k], v);
    x[i] = i >= lr ? v : 0.0;
  }
  double myd = dd[0], myinv = di[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) { myd = lr == i ? dd[i] : myd; myinv = lr == i ? di[i] : myinv; }
  __syncwarp();  // every lane has read the block (and, J > 0, the tiles of the gather)
  if (lane < 8) {
    sm[L::O_D + 8 * J + lr] = myd; sm[L::O_INVD + 8 * J + lr] = myinv;
    double* __restrict__ tile = T + L::tile(J, J);
#pragma unroll
    for (int i = 0; i < 8; ++i) tile[wd_el(i, lr)] = x[i];  // S[i][lr]
  }
  __syncwarp();
  // ---- panel solve L(I,J) = C(I,J) S' D^-1 from the accumulators, look-ahead update of the next diagonal tile ----
  if constexpr (CNT1 > 0) {
    const double idr = sm[L::O_INVD + 8 * J + r];
    const double bt0 = T[L::tile(J, J) + lane] * idr, bt1 = T[L::tile(J, J) + hi] * idr;
#pragma unroll
    for (int i = 0; i < CNT1; ++i) {
      double a0, a1, p0 = 0.0, p1 = 0.0, q0 = 0.0, q1 = 0.0;
      wdl_c2a(acc[i][0], acc[i][1], lane, a0, a1);
      dmma884(p0, p1, a0, bt0);
      dmma884(q0, q1, a1, bt1);
      p0 += q0; p1 += q1;
      *reinterpret_cast<double2*>(T + L::tile(J + 1 + i, J) + coff) = make_double2(p0, p1);
      if (HASD && i == 0) {
        double l0, l1;
        wdl_c2a(p0, p1, lane, l0, l1);
        const double bd0 = l0 * -sm[L::O_D + 8 * J + q], bd1 = l1 * -sm[L::O_D + 8 * J + 4 + q];
        double e0 = 0.0, e1 = 0.0;
        dmma884(accD[0], accD[1], l0, bd0);
        dmma884(e0, e1, l1, bd1);
        *reinterpret_cast<double2*>(T + L::tile(J + 1, J + 1) + coff) = make_double2(accD[0] + e0, accD[1] + e1);
      }
    }
  }
  // ---- forward substitution, second half: w_J = D_J^-1 S_J (b_J - sum) ----
  if constexpr (FWD) {
    double s = (fsa + fsb) + (fsc + fsd);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    const double tr = sm[L::O_R + 8 * J + r] - s;  // t[r] (replicated over q)
    const int toff = wd_el(q, r);                  // element [q][r]; [4 + q][r] is 16 further
    const double* __restrict__ S = T + L::tile(J, J);
    double z0 = S[toff] * tr, z1 = S[toff + 16] * tr;  // S[4h + q][r] t[r]
    z0 = wd_sum_r(z0); z1 = wd_sum_r(z1);
    if (r == 0) {
      sm[L::O_W + 8 * J + q] = z0 * sm[L::O_INVD + 8 * J + q];
      sm[L::O_W + 8 * J + 4 + q] = z1 * sm[L::O_INVD + 8 * J + 4 + q];
    }
  }
  __syncwarp();
}

template <int N, int K, int J>
__device__ __forceinline__ void wdl_columns(double* __restrict__ sm, int lane, double& chk) {
  if constexpr (J < WdLayout<N, K>::NT) {
    wdl_column<N, K, J>(sm, lane, chk);
    wdl_columns<N, K, J + 1>(sm, lane, chk);
  }
}

template <int N, int K, int NW, int J, int LOCK = 1>
__device__ __forceinline__ void wd_columns(double* __restrict__ sm, int lane, int warp, double& chk) {
  if constexpr (J < WdLayout<N, K>::NT) {
    if constexpr (NW == 1) wd_column<N, K, J, LOCK>(sm, lane, chk);
    else wd2_column<N, K, J>(sm, lane, warp, chk);
    wd_columns<N, K, NW, J + 1, LOCK>(sm, lane, warp, chk);
  }
}

// Factor the KKT matrix held in the tiles and solve for the right-hand side held in row 0 of tile row NT.
// On return sm[O_XV .. O_XV + P) = K^-1 rhs and the cp.async staging of G, A_eq for the next trip is in flight. Returns false on a zero or non-finite pivot (-> MO_QP_FAILED_FACTORIZATION,
// the batched stand-in for the exception of qp.cc:303-307); with two warps only warp 0 holds the answer.
template <int N, int K, int NW, bool LA, bool PBS, int LOCK = 1>
__device__ __forceinline__ bool wd_factor_solve(double* __restrict__ sm, int lane, int warp,
                                                const double* __restrict__ Gp, const double* __restrict__ Ap) {
  using L = WdLayout<N, K>;
  constexpr bool kStageInBacksub = WdLayout<N, K>::STAGE_IN_BACKSUB;
  double chk = 0.0;
  if constexpr (LOCK > 1) __syncthreads_or(1);  // the factorizations of the CTA's warps start together (see the kernel's tail)
  WD_T(1);
  if constexpr (LA) wdl_columns<N, K, 0>(sm, lane, chk);
  else wd_columns<N, K, NW, 0, LOCK>(sm, lane, warp, chk);
  if constexpr (LA) WD_T(3);
  if (NW == 2 && warp != 0) return true;
  // back substitution L' x = w, w = D^-1 L^-1 rhs = row 0 of tile row NT; block J: t = w_J - sum_{I > J} L(I,J)' x_I,
  // x_J = inv(L_JJ)' t with S = inv(L_JJ) stored in the diagonal tile.
  const double* __restrict__ T = sm;
  double* __restrict__ XS = sm + L::O_XV;
  const int r = lane >> 2, q = lane & 3;
  const int toff = wd_el(q, r);  // element [q][r]; [4 + q][r] is 16 further
  if constexpr (PBS) {
    // Software-pipelined and fully unrolled: the only term of block J on the dependent chain is the one of the block
    // just solved, x_{J+1}, which arrives REPLICATED in registers (wd_sum_r leaves x_J[q], x_J[4 + q] in every lane with
    // that q -- the operand layout of the next product, so no store / barrier / load sits on the chain); the terms of
    // the older blocks (I >= J + 2) were summed one block earlier, under the shuffles of that block's chain.
    double xr0 = 0.0, xr1 = 0.0, pa = 0.0, pb = 0.0;
#pragma unroll
    for (int JJ = 0; JJ < L::NT; ++JJ) {
      const int J = L::NT - 1 - JJ;
      const double* __restrict__ tj = T + 64 * (J * L::NTR - (J * (J - 1)) / 2);  // tile (J, J); tile (I, J) = tj + 64 (I - J)
      double sa = pa, sb = pb;
      if (J + 1 < L::NT) { sa = fma(tj[64 + toff], xr0, sa); sb = fma(tj[64 + toff + 16], xr1, sb); }
      double sq = sa + sb;
      sq += __shfl_xor_sync(0xffffffffu, sq, 1);
      sq += __shfl_xor_sync(0xffffffffu, sq, 2);
      const double wr = L::RHS_ROW ? tj[64 * (L::NT - J) + wd_el(0, r)] : sm[L::O_W + 8 * J + r];
      const double tr = wr - sq;  // t[r] (replicated over q)
      double x0 = tj[lane] * tr, x1 = tj[32 + (lane ^ 8)] * tr;  // S[r][4h + q] t[r]
      x0 = wd_sum_r(x0); x1 = wd_sum_r(x1);
      // partial sums of block J - 1 over I >= J + 1 (x_I in shared memory since the previous trip's barrier)
      double na = 0.0, nb = 0.0, nc = 0.0, nd = 0.0;
      if (J > 0) {
        const double* __restrict__ tp = T + 64 * ((J - 1) * L::NTR - ((J - 1) * (J - 2)) / 2);  // tile (J-1, J-1)
#pragma unroll
        for (int I = J + 1; I < L::NT; I += 2) {
          const double* __restrict__ ti = tp + 64 * (I - J + 1);
          na = fma(ti[toff], XS[8 * I + q], na);
          nb = fma(ti[toff + 16], XS[8 * I + 4 + q], nb);
          if (I + 1 < L::NT) {
            nc = fma(ti[64 + toff], XS[8 * I + 8 + q], nc);
            nd = fma(ti[64 + toff + 16], XS[8 * I + 12 + q], nd);
          }
        }
      }
      pa = na + nc; pb = nb + nd;
      xr0 = x0; xr1 = x1;
      if (r == 0) { XS[8 * J + q] = x0; XS[8 * J + 4 + q] = x1; }
      __syncwarp();
      if (kStageInBacksub && J < L::NTN) {  // block column J of L is consumed: stage the next trip's G, A into it
        double* __restrict__ dst = sm + 64 * (J * L::NTR - (J * (J - 1)) / 2) + lane;
        const int hi = 32 + (lane ^ 8) - lane;
        const double* __restrict__ g = Gp + (q + 8 * J) * N + r + 8 * J;
#pragma unroll
        for (int I = J; I < L::NTN; ++I, dst += 64, g += 8) { cp_async8(dst, g); cp_async8(dst + hi, g + 4 * N); }
        const double* __restrict__ ap = Ap + (q + 8 * J) * K + r;
#pragma unroll
        for (int I = L::NTN; I < L::NT; ++I, dst += 64, ap += 8) { cp_async8(dst, ap); cp_async8(dst + hi, ap + 4 * K); }
      }
    }
  } else {
#pragma unroll 1
  for (int J = L::NT - 1; J >= 0; --J) {
    const double* __restrict__ tj = T + 64 * (J * L::NTR - (J * (J - 1)) / 2);  // tile (J, J); tile (I, J) = tj + 64 (I - J)
    // t[r] = w[r] - sum_I sum_k L(I,J)[k][r] x_I[k]: the tiles are read TRANSPOSED (element [4h + q][r], two contiguous
    // 128-byte runs per request in this layout), so the sum over k is two accumulators per lane plus a reduction over
    // q (2 shuffle levels), and t lands where the second product needs it (lane group r) without a broadcast.
    double sa = 0.0, sb = 0.0, sc = 0.0, sd = 0.0;
    for (int I = J + 1; I < L::NT; I += 2) {
      const double* __restrict__ ti = tj + 64 * (I - J);
      sa = fma(ti[toff], XS[8 * I + q], sa);
      sb = fma(ti[toff + 16], XS[8 * I + 4 + q], sb);
      if (I + 1 < L::NT) {
        sc = fma(ti[64 + toff], XS[8 * I + 8 + q], sc);
        sd = fma(ti[64 + toff + 16], XS[8 * I + 12 + q], sd);
      }
    }
    double sq = (sa + sb) + (sc + sd);
    sq += __shfl_xor_sync(0xffffffffu, sq, 1);
    sq += __shfl_xor_sync(0xffffffffu, sq, 2);
    const double wr = L::RHS_ROW ? tj[64 * (L::NT - J) + wd_el(0, r)] : sm[L::O_W + 8 * J + r];
    const double tr = wr - sq;                           // t[r] (replicated over q)
    double x0 = tj[lane] * tr, x1 = tj[32 + (lane ^ 8)] * tr;  // S[r][4h + q] t[r]
    x0 = wd_sum_r(x0); x1 = wd_sum_r(x1);
    if (r == 0) { XS[8 * J + q] = x0; XS[8 * J + 4 + q] = x1; }
    __syncwarp();
    // block column J of L is consumed: stage the next trip's G, A into it now, under the rest of the substitution
    if (kStageInBacksub && J < L::NTN) {
      double* __restrict__ dst = sm + 64 * (J * L::NTR - (J * (J - 1)) / 2) + lane;
      const int hi = 32 + (lane ^ 8) - lane;  // second half of the tile, swizzled
      const double* __restrict__ g = Gp + (q + 8 * J) * N + r + 8 * J;
      for (int I = J; I < L::NTN; ++I, dst += 64, g += 8) {
        cp_async8(dst, g);
        cp_async8(dst + hi, g + 4 * N);
      }
      const double* __restrict__ ap = Ap + (q + 8 * J) * K + r;
      for (int I = L::NTN; I < L::NT; ++I, dst += 64, ap += 8) {
        cp_async8(dst, ap);
        cp_async8(dst + hi, ap + 4 * K);
      }
    }
  }
  }
  if (kStageInBacksub) cp_async_commit();
  WD_T(5);
  return chk == chk;
}

// ---------------------------------------------------------------------------------------------------------
// The solve kernel: persistent CTAs of NW warps, one problem per CTA at a time. NW = 1: everything in one warp.
// NW = 2: warp 0 owns the iteration (vectors, bookkeeping, diagonal blocks, panel solves, back substitution), warp 1
// is the look-ahead helper (gather of the next block column, half of the staging, the A_eq rows of the matvec).
// Lane l of warp 0 carries variables v = l + 32 i (x, both slacks / multipliers / bounds, c) and equality row l.
// ---------------------------------------------------------------------------------------------------------
template <int NW>
__device__ __forceinline__ void wd_bar() {
  if (NW == 1) __syncwarp(); else __syncthreads();
}

template <int N, int K>
__host__ __device__ constexpr size_t wd_region_doubles() { return (WdLayout<N, K>::TOTAL + 1) & ~size_t(1); }  // per warp, 16-byte multiple

template <int N, int K, int NW, int MINB, bool LA, bool PBS, int LOCK = 1>
__global__ void __launch_bounds__(32 * NW * LOCK, MINB) qp_wd_kernel(const QPArgs a) {
  static_assert(!LA || NW == 1, "the look-ahead schedule is a one-warp schedule");
  static_assert(LOCK == 1 || (NW == 1 && !LA), "lock-step CTAs: one warp per problem, plain columns");
  using L = WdLayout<N, K>;
  constexpr int P = L::P, M = L::M, NT = L::NT, NTN = L::NTN, VPL = L::VPL, STRIDE = N + 2 * M + K;
  static_assert(NW == 1 || NW == 2, "one or two warps per problem");
  extern __shared__ __align__(16) double sm_base[];
  double* const sm = sm_base + (LOCK > 1 ? (threadIdx.x >> 5) * wd_region_doubles<N, K>() : 0);
  const int lane = threadIdx.x & 31, warp = NW == 1 ? 0 : (int)(threadIdx.x >> 5), r = lane >> 2, q = lane & 3;
  const bool main_warp = warp == 0, helper = warp == NW - 1;  // NW = 1: the one warp plays both roles
  double* __restrict__ T = sm;
  double* __restrict__ XV = sm + L::O_XV;
  double* __restrict__ R = sm + L::O_R;
  double* __restrict__ RC = sm + L::O_RC;
  volatile int* ctl = reinterpret_cast<volatile int*>(sm + L::O_CTL);
  const double tol2 = a.kkt_tol * a.kkt_tol, tau = 0.995;  // qp.cc:192
  const bool isy = lane < K;
#ifdef MO_WD_TRACE
  if (threadIdx.x == 0) { for (int i = 0; i < 16; ++i) s_wd_tr[i] = 0; s_wd_t0 = clock64(); }
  wd_bar<NW>();
#endif

  for (long long prob = LOCK > 1 ? (long long)blockIdx.x * LOCK + (threadIdx.x >> 5) : (long long)blockIdx.x; prob < a.batch;
       prob += (long long)gridDim.x * LOCK) {
    const double* __restrict__ Gp = a.G + prob * (N * N);
    const double* __restrict__ Ap = a.A + prob * (N * K);
    double* __restrict__ vp = a.vars + prob * STRIDE;
    wd_bar<NW>();
    wd_stage<N, K, NW>(T, Gp, Ap, lane, warp);
    // right-hand-side tile row: rows 1..7 stay zero for the life of the problem
    if (helper && L::RHS_ROW) {
#pragma unroll
      for (int J = 0; J < NT; ++J) { T[L::tile(NT, J) + lane] = 0.0; T[L::tile(NT, J) + 32 + (lane ^ 8)] = 0.0; }
    }
    double x[VPL], s0[VPL], s1[VPL], z0[VPL], z1[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) { x[i] = 0.0; s0[i] = s1[i] = z0[i] = z1[i] = 1.0; }
    // c, b_eq and the box bounds are NOT held in registers for the life of the problem: the three places that need
    // them re-read their 40 bytes per lane (L1 / L2 hits), and the factorization in between gets the registers
    const double* __restrict__ cp = a.c + prob * N;
    const double* __restrict__ ibp = a.ib + prob * M;
    auto load_bounds = [&](double (&b0)[VPL], double (&b1)[VPL]) {
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        const int v = lane + 32 * i;
        b0[i] = b1[i] = 1.0;
        if (v < N && main_warp) {
          const double2 bb = __ldg(reinterpret_cast<const double2*>(ibp + 2 * v));
          b0[i] = bb.x; b1[i] = bb.y;
        }
      }
    };
    double y = 0.0;
    int status = -1, iters = 0;
    // ComputeInitialGuess tail for my variables (qp.cc:462-481): clamp in list order (ClampX, qp.hpp:43-53),
    // s = max(1e-9, a x + b), z = 1/s.
    auto clamp_init = [&]() {
      double b0[VPL], b1[VPL];
      load_bounds(b0, b1);
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        double xv = x[i];
        xv = xv > -b0[i] ? xv : -b0[i];
        xv = xv < b1[i] ? xv : b1[i];
        x[i] = xv;
        s0[i] = fmax(1.0e-9, xv + b0[i]); s1[i] = fmax(1.0e-9, b1[i] - xv);
        z0[i] = 1.0 / s0[i]; z1[i] = 1.0 / s1[i];
      }
    };
    bool guess_pending = a.guess == MO_GUESS_SOLVE_EQUALITY_CONSTRAINED;  // uniform over the CTA
    if (a.guess == MO_GUESS_USER_PROVIDED) {
      if (main_warp) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
          const int v = lane + 32 * i;
          if (v < N) {
            x[i] = vp[v];
            s0[i] = vp[N + 2 * v]; s1[i] = vp[N + 2 * v + 1];
            z0[i] = vp[N + M + K + 2 * v]; z1[i] = vp[N + M + K + 2 * v + 1];
          }
        }
        if (isy) y = vp[N + M + lane];
      }
    } else if (!guess_pending) {
      clamp_init();
    }
    double mu = a.initial_mu, alpha_p = 1.0, alpha_d = 1.0;
    double kkt[4] = {0.0, 0.0, 0.0, 0.0};
    bool first = true;
    // ---- Solve loop (qp.cc:110-149); each trip = residuals of the current state, bookkeeping, one Newton step ----
    for (;;) {
      double rhsv[VPL], rhse = 0.0;
      // publish [x; -y] for the matvec
      if (main_warp) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) if (lane + 32 * i < N) XV[lane + 32 * i] = x[i];
        if (isy) XV[N + lane] = -y;
      }
      cp_async_wait_all();
      wd_bar<NW>();
      // the K x K zero block (the factorization overwrote it)
      if (helper) {
#pragma unroll
        for (int J = NTN; J < NT; ++J)
#pragma unroll
          for (int I = J; I < NT; ++I) { T[L::tile(I, J) + lane] = 0.0; T[L::tile(I, J) + 32 + (lane ^ 8)] = 0.0; }
      }
      const double be = (isy && main_warp) ? __ldg(a.b + prob * K + lane) : 0.0;
      if (guess_pending) {  // KKT system without Sigma, rhs = -[c; b_eq]
#pragma unroll
        for (int i = 0; i < VPL; ++i) rhsv[i] = (lane + 32 * i < N && main_warp) ? -__ldg(cp + lane + 32 * i) : 0.0;
        rhse = -be;
      } else {
        // (1) EvaluateKKTConditions (qp.cc:391-420): symmetric matvec [[G, A'],[A, 0]] [x; -y] over the stored tiles;
        //     the main warp takes the G tile rows, the helper the A_eq tile rows
        double colacc[NTN][2];
#pragma unroll
        for (int J = 0; J < NTN; ++J) colacc[J][0] = colacc[J][1] = 0.0;
#pragma unroll
        for (int I = 0; I < NT; ++I) {
          if (I < NTN ? !main_warp : !helper) continue;
          double ra = 0.0, rb = 0.0;
          const double vr = XV[8 * I + r];
#pragma unroll
          for (int J = 0; J <= (I < NTN ? I : NTN - 1); ++J) {
            const double e0 = T[L::tile(I, J) + lane], e1 = T[L::tile(I, J) + 32 + (lane ^ 8)];  // [r][q], [r][4 + q]
            const double v0 = XV[8 * J + q], v1 = XV[8 * J + 4 + q];
            if (I == J) {  // diagonal tile: lower triangle significant
              ra = q <= r ? fma(e0, v0, ra) : ra;
              rb = 4 + q <= r ? fma(e1, v1, rb) : rb;
              colacc[J][0] = q < r ? fma(e0, vr, colacc[J][0]) : colacc[J][0];
              colacc[J][1] = 4 + q < r ? fma(e1, vr, colacc[J][1]) : colacc[J][1];
            } else {
              ra = fma(e0, v0, ra); rb = fma(e1, v1, rb);
              colacc[J][0] = fma(e0, vr, colacc[J][0]);
              colacc[J][1] = fma(e1, vr, colacc[J][1]);
            }
          }
          double rs = ra + rb;
          rs += __shfl_xor_sync(0xffffffffu, rs, 1);
          rs += __shfl_xor_sync(0xffffffffu, rs, 2);
          if (q == 0) R[8 * I + r] = rs;
        }
        __syncwarp();
#pragma unroll
        for (int J = 0; J < NTN; ++J) {
          const double c0 = wd_sum_r(colacc[J][0]), c1 = wd_sum_r(colacc[J][1]);
          if (r == 0) {
            if (NW == 1) { R[8 * J + q] += c0; R[8 * J + 4 + q] += c1; }
            else if (main_warp) { R[8 * J + q] += c0; R[8 * J + 4 + q] += c1; }
            else { RC[8 * J + q] = c0; RC[8 * J + 4 + q] = c1; }
          }
        }
        wd_bar<NW>();
        WD_T(0);
        if (main_warp) {
        // (2) per-variable residuals, Sigma, r_aug pieces; norms (ComputeErrors, qp.cc:423-437)
        double red[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // |r_d|^2, |r_pe|^2, |r_pi|^2, sum (sz)^2, sum sz, #(s <= 0)
        // p (r_pi), 1/s and w = z/s are NOT kept across the factorization: step (5) recomputes them from the unchanged
        // x, s, z, bounds (bit-identical, ~12 FP64 operations per variable) and the factorization gets the registers
        double t1[VPL], t2[VPL], p0[VPL], p1[VPL], w0[VPL], w1[VPL], i0[VPL], i1[VPL], b0[VPL], b1[VPL];
        load_bounds(b0, b1);
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
          const int v = lane + 32 * i;
          t1[i] = t2[i] = p0[i] = p1[i] = w0[i] = w1[i] = i0[i] = i1[i] = 0.0;
          if (v < N) {
            const double base = (NW == 1 ? R[v] : R[v] + RC[v]) + __ldg(cp + v);
            p0[i] = (x[i] + b0[i]) - s0[i]; p1[i] = (b1[i] - x[i]) - s1[i];
            i0[i] = fast_rcp(s0[i]); i1[i] = fast_rcp(s1[i]);
            w0[i] = z0[i] * i0[i]; w1[i] = z1[i] * i1[i];
            const double rdv = (base - z0[i]) + z1[i];
            const double c0 = s0[i] * z0[i], c1 = s1[i] * z1[i];
            red[0] += rdv * rdv;
            red[2] += fma(p0[i], p0[i], p1[i] * p1[i]);
            red[3] += fma(c0, c0, c1 * c1);
            red[4] += c0 + c1;
            red[5] += ((s0[i] > 0.0) ? 0.0 : 1.0) + ((s1[i] > 0.0) ? 0.0 : 1.0);
            t1[i] = base + fma(w0[i], p0[i], -(w1[i] * p1[i]));  // r_aug without the mu term (qp.cc:337-342)
            t2[i] = i0[i] - i1[i];
          }
        }
        double pe = 0.0;
        if (isy) { pe = R[N + lane] + be; red[1] = pe * pe; }
#pragma unroll
        for (int i = 0; i < 6; ++i) red[i] = wd_sum(red[i]);
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
          if (a.iter_log && lane == 0) {
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
        if (status < 0) {
          if (a.iter_log && lane == 0) {  // kkt_initial of the coming iteration (ComputeErrors with the new mu, qp.cc:118)
            double* row = a.iter_log + (prob * a.max_iterations + iters) * MO_QP_ITER_LOG_STRIDE;
            const double cr = red[3] - 2 * (red[4] * mu) + (mu * mu) * M;
            row[0] = sqrt(red[0]); row[1] = sqrt(cr > 0.0 ? cr : 0.0); row[2] = sqrt(red[1]); row[3] = sqrt(red[2]);
          }
          // (4) ComputeLDLT assembly (qp.cc:289-297): Sigma on the diagonal; rhs = -[r_aug; r_pe]
#pragma unroll
          for (int i = 0; i < VPL; ++i) {
            const int v = lane + 32 * i;
            rhsv[i] = -fma(-mu, t2[i], t1[i]);
            if (v < N) {
              const int jv = v >> 3, rr = v & 7;
              T[64 * (jv * L::NTR - (jv * (jv - 1)) / 2) + wd_el(rr, rr)] += w0[i] + w1[i];
            }
          }
          rhse = -pe;
        }
        }  // main_warp
      }
      // right-hand side into row 0 of tile row NT, or into the vector the block forward substitution reads
      if (main_warp && status < 0 && !L::RHS_ROW) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) if (lane + 32 * i < N) R[lane + 32 * i] = rhsv[i];
        if (isy) R[N + lane] = rhse;
      }
      if (main_warp && status < 0 && L::RHS_ROW) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
          const int v = lane + 32 * i;
          if (v < N) {
            const int jv = v >> 3, cc = v & 7;
            T[64 * (jv * L::NTR - (jv * (jv - 1)) / 2 + (NT - jv)) + wd_el(0, cc)] = rhsv[i];
          }
        }
        if (isy) {
          const int e = N + lane, je = e >> 3, cc = e & 7;
          T[64 * (je * L::NTR - (je * (je - 1)) / 2 + (NT - je)) + wd_el(0, cc)] = rhse;
        }
      }
      if (NW == 2) {  // the helper learns whether the loop goes on
        if (threadIdx.x == 0) *ctl = status;
        __syncthreads();
        status = *ctl;
      } else {
        __syncwarp();
      }
      if (status >= 0) break;
      bool ok = wd_factor_solve<N, K, NW, LA, PBS, LOCK>(sm, lane, warp, Gp, Ap);
      if (NW == 2) {
        if (threadIdx.x == 0) *ctl = ok ? 1 : 0;
        __syncthreads();  // also: the back substitution is done, the tiles are free for the next trip's staging
        ok = *ctl != 0;
      }
      if (!ok) { status = MO_QP_FAILED_FACTORIZATION; break; }
      if (!L::STAGE_IN_BACKSUB) wd_stage<N, K, NW>(T, Gp, Ap, lane, warp);  // next trip's G, A: overlaps the update below
      WD_T(6);
      if (guess_pending) {  // x <- dx, y <- dy of the equality constrained solve (qp.cc:452-459), then clamp / s / z
        guess_pending = false;
        if (main_warp) {
#pragma unroll
          for (int i = 0; i < VPL; ++i) if (lane + 32 * i < N) x[i] = XV[lane + 32 * i];
          if (isy) y = -XV[N + lane];
          clamp_init();
        }
        wd_bar<NW>();  // XV consumed before the next trip republishes it
        continue;
      }
      if (main_warp) {
      // (5) ds, dz (qp.cc:359-363), step lengths (:485-507), state update (:196-199)
      double al0 = 1.0, al1 = 1.0, dx[VPL], d0[VPL], d1[VPL], e0[VPL], e1[VPL], b0[VPL], b1[VPL];
      load_bounds(b0, b1);
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        const int v = lane + 32 * i;
        dx[i] = d0[i] = d1[i] = e0[i] = e1[i] = 0.0;
        if (v < N) {
          const double p0 = (x[i] + b0[i]) - s0[i], p1 = (b1[i] - x[i]) - s1[i];  // as in step (2)
          const double i0 = fast_rcp(s0[i]), i1 = fast_rcp(s1[i]);
          const double w0 = z0[i] * i0, w1 = z1[i] * i1;
          dx[i] = XV[v];
          d0[i] = dx[i] + p0; d1[i] = p1 - dx[i];
          e0[i] = fma(mu, i0, -fma(w0, d0[i], z0[i]));
          e1[i] = fma(mu, i1, -fma(w1, d1[i], z1[i]));
          if (s0[i] + d0[i] <= 0.0 && fabs(d0[i]) > 0) al0 = fmin(al0, -tau * s0[i] / d0[i]);
          if (s1[i] + d1[i] <= 0.0 && fabs(d1[i]) > 0) al0 = fmin(al0, -tau * s1[i] / d1[i]);
          if (z0[i] + e0[i] <= 0.0 && fabs(e0[i]) > 0) al1 = fmin(al1, -tau * z0[i] / e0[i]);
          if (z1[i] + e1[i] <= 0.0 && fabs(e1[i]) > 0) al1 = fmin(al1, -tau * z1[i] / e1[i]);
        }
      }
      alpha_p = wd_min(al0); alpha_d = wd_min(al1);
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        x[i] = fma(dx[i], alpha_p, x[i]);
        s0[i] = fma(d0[i], alpha_p, s0[i]); s1[i] = fma(d1[i], alpha_p, s1[i]);
        z0[i] = fma(e0[i], alpha_d, z0[i]); z1[i] = fma(e1[i], alpha_d, z1[i]);
      }
      if (isy) y = fma(-XV[N + lane], alpha_d, y);
      }
      WD_T(7);
    }
    // ---- results ----
    cp_async_wait_all();
    if (main_warp) {
#pragma unroll
      for (int i = 0; i < VPL; ++i) {
        const int v = lane + 32 * i;
        if (v < N) {
          vp[v] = x[i];
          *reinterpret_cast<double2*>(vp + N + 2 * v) = make_double2(s0[i], s1[i]);
          *reinterpret_cast<double2*>(vp + N + M + K + 2 * v) = make_double2(z0[i], z1[i]);
        }
      }
      if (isy) vp[N + M + lane] = y;
      if (a.lagrange) {  // ComputeLagrangeMultiplierSummary, qp.cc:539-546
        const double mn = wd_min(isy ? y : 1.0e300), li = -wd_min(isy ? -fabs(y) : 0.0);
        if (lane == 0) { a.lagrange[prob * 2] = mn; a.lagrange[prob * 2 + 1] = li; }
      }
      if (lane == 0) {
        a.status[prob] = status;
        a.iters[prob] = iters;
        if (a.kkt_final) {
          double* kf = a.kkt_final + prob * 4;
          kf[0] = kkt[0]; kf[1] = kkt[1]; kf[2] = kkt[2]; kf[3] = kkt[3];
        }
      }
    }
    WD_T(8);
  }
  if constexpr (LOCK > 1) {
    // out of problems: keep the barrier count of the warps that still factor (one __syncthreads_or + wd_lock_barriers per
    // factorization) until every warp of the CTA votes "done"
    while (__syncthreads_or(0)) {
#pragma unroll 1
      for (int i = 0; i < wd_lock_barriers<N, K>(); ++i) __syncthreads();
    }
  }
#ifdef MO_WD_TRACE
  if (blockIdx.x == 0 && threadIdx.x == 0)
    printf("wd trace n=%d: matvec %lld | pervar+bookkeeping+rhs %lld | gather %lld | diag+inv %lld | panel %lld | backsub %lld | stage %lld | update %lld | results+load %lld\n",
           N, s_wd_tr[0], s_wd_tr[1], s_wd_tr[2], s_wd_tr[3], s_wd_tr[4], s_wd_tr[5], s_wd_tr[6], s_wd_tr[7], s_wd_tr[8]);
#endif
}

}  // namespace mo
