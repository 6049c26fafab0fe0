// Gauss-Newton / Levenberg-Marquardt linearization for MEDIUM dense problems (n = 32 .. 64, a multiple of 8) on the FP64
// tensor path: G = J'J (+ lambda I), c = J'r with mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4), one WARP per problem.
//
// Replaces, per problem: ConstrainedNonlinearLeastSquares::LinearizeAndFillQP (source/nonlinear.cc:170-214) with
// Residual::Model::UpdateHessian (include/mini_opt/residual.hpp:186-226: the O(R P^2) double loop over
// J.col(i).dot(J.col(j)), lower triangle only, :216-220) and UpdateJacobian (:230-250), the LM damping of
// nonlinear.cc:187-189 and the ShiftTo of the inequality constraints (:209-212, qp.hpp:57-59).
//
// Why DMMA here and DFMA below n = 32 (north_star item 4): the contraction J'J is N x R x N; at N >= 32 every 8x8 output
// tile consumes R / 4 full 8x8x4 products, i.e. the tensor path does 256 FMA per issued instruction against 32 for DFMA,
// and each loaded element of J feeds up to 2 N / 8 products from registers.
//
// Mapping. Output tile (I, J), I >= J, of the lower block triangle lives in one DMMA accumulator fragment for the whole
// contraction (NT (NT + 1) / 2 fragments = 72 doubles per lane at n = 64): nothing is staged in shared memory, J is read
// ONCE from global memory, straight into the operand fragments. For rows 8g .. 8g+7 of J and tile column T, lane
// (r, q) loads the two consecutive rows 8g + 2q, 8g + 2q + 1 of column 8T + r with one 16-byte request (64 contiguous
// bytes per column and warp request). The two halves of that register pair are the operands of two k-steps (k-sets
// {0,2,4,6} and {1,3,5,7} of the row group): the SAME register is the A fragment of tile row T (A[m = r][k = q] =
// J[k][8T + r]) and the B fragment of tile column T (B[k = q][n = r]), so every element is loaded once and used
// 2 NT times. The fragments of row group g + 1 are requested before the products of group g are issued.
// J'r rides along on DFMA (one fused multiply-add per loaded element, reduced over q at the end), f = 1/2 |r|^2 too.
// The result is written in the layout the QP kernels read: G column-major n x n, lower triangle, strict upper triangle
// exactly 0 (test/residual_test.cc:129-133), lambda added on the diagonal only if lambda > 0 (nonlinear.cc:187).
//
// Roofline: R N (N + 1) + 2 R N + N algorithmic flops against 8 (R N + R + N^2 + N) bytes: 5.4 flop/B at
// N = 64, R = 128, within 10 % of the machine balance (37 TFLOP/s DMMA / 6.5 TB/s HBM): the kernel is bound by both.
#pragma once
#include "qp_dmma.cuh"
#include "nls_chain.cuh"

namespace mo {

template <int NT>
__global__ void __launch_bounds__(32) nls_linearize_dmma_kernel(const LinearizeArgs a) {
  constexpr int N = 8 * NT, NACC = NT * (NT + 1) / 2;
  const int lane = threadIdx.x, r = lane >> 2, q = lane & 3;
  const int R = a.r_cost, k = a.r_eq, m = a.m;
  const int groups = (R + 7) / 8;
  for (long long p = blockIdx.x; p < a.batch; p += gridDim.x) {
    const double* __restrict__ Jp = a.J_cost + p * (long long)R * N;
    const double* __restrict__ rp = a.r_cost_v + p * R;
    double acc[NACC][2];
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i][0] = acc[i][1] = 0.0;
    double cj[NT];  // J'r: partial sums of column 8T + r over this lane's rows
#pragma unroll
    for (int t = 0; t < NT; ++t) cj[t] = 0.0;
    double fsum = 0.0;
    // fragment pair of row group g: rows 8g + 2q, 8g + 2q + 1 of column 8T + r (zero beyond R)
    auto load_group = [&](int g, double2 (&f)[NT], double2& rr) {
      const int row = 8 * g + 2 * q;
      const bool in = row + 1 < R, half = row < R;
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        const double* src = Jp + (long long)(8 * t + r) * R + row;
        if (in) f[t] = __ldg(reinterpret_cast<const double2*>(src));
        else f[t] = make_double2(half ? __ldg(src) : 0.0, 0.0);
      }
      if (in) rr = __ldg(reinterpret_cast<const double2*>(rp + row));
      else rr = make_double2(half ? __ldg(rp + row) : 0.0, 0.0);
    };
    double2 cur[NT], nxt[NT], rcur, rnxt;
    load_group(0, cur, rcur);
    for (int g = 0; g < groups; ++g) {
      if (g + 1 < groups) load_group(g + 1, nxt, rnxt);
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        cj[t] = fma(cur[t].x, rcur.x, cj[t]);
        cj[t] = fma(cur[t].y, rcur.y, cj[t]);
      }
      if (r == 0) fsum = fma(rcur.x, rcur.x, fma(rcur.y, rcur.y, fsum));
      int idx = 0;
#pragma unroll
      for (int J = 0; J < NT; ++J) {
#pragma unroll
        for (int I = J; I < NT; ++I, ++idx) {
          dmma884(acc[idx][0], acc[idx][1], cur[I].x, cur[J].x);
          dmma884(acc[idx][0], acc[idx][1], cur[I].y, cur[J].y);
        }
      }
#pragma unroll
      for (int t = 0; t < NT; ++t) cur[t] = nxt[t];
      rcur = rnxt;
    }
    // ---- outputs ----
    const double lambda = a.lambda[p];
    double* __restrict__ Gp = a.G + p * (long long)N * N;
    int idx = 0;
#pragma unroll
    for (int J = 0; J < NT; ++J) {
      // strict upper block triangle of this block column: exactly zero
#pragma unroll
      for (int I = 0; I < J; ++I) {
        Gp[(long long)(8 * J + 2 * q) * N + 8 * I + r] = 0.0;
        Gp[(long long)(8 * J + 2 * q + 1) * N + 8 * I + r] = 0.0;
      }
#pragma unroll
      for (int I = J; I < NT; ++I, ++idx) {
        const int row = 8 * I + r, c0 = 8 * J + 2 * q, c1 = c0 + 1;
        double v0 = acc[idx][0], v1 = acc[idx][1];
        if (I == J) {
          if (row == c0 && lambda > 0) v0 += lambda;  // nonlinear.cc:187-189
          if (row == c1 && lambda > 0) v1 += lambda;
          if (row < c0) v0 = 0.0;
          if (row < c1) v1 = 0.0;
        }
        Gp[(long long)c0 * N + row] = v0;
        Gp[(long long)c1 * N + row] = v1;
      }
    }
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      double v = cj[t];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      if (q == 0) a.c[p * N + 8 * t + r] = v;
    }
    fsum += __shfl_xor_sync(0xffffffffu, fsum, 1);
    fsum += __shfl_xor_sync(0xffffffffu, fsum, 2);
    // equality rows, shifted bounds, Errors{f, equality}
    const double* __restrict__ xp = a.x + p * N;
    double eq = 0.0;
    if (k > 0) {
      const double* __restrict__ Je = a.J_eq + p * (long long)k * N;
      double* __restrict__ Ap = a.A_eq + p * (long long)k * N;
      for (int e = lane; e < k * N; e += 32) Ap[e] = __ldg(Je + e);
      for (int e = lane; e < k; e += 32) {
        const double v = __ldg(a.r_eq_v + p * k + e);
        a.b_eq[p * k + e] = v;
        eq += fabs(v);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) eq += __shfl_xor_sync(0xffffffffu, eq, o);
    }
    for (int i = lane; i < m; i += 32)
      a.ineq_b_shifted[p * m + i] = a.ineq_a[p * m + i] * xp[a.ineq_var[p * m + i]] + a.ineq_b[p * m + i];  // qp.hpp:57-59
    if (lane == 0) {
      a.errors[2 * p] = 0.5 * fsum;
      a.errors[2 * p + 1] = eq;
    }
  }
}

}  // namespace mo
