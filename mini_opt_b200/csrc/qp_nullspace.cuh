// Batched QPNullSpaceSolver (equality-constrained QP by the null-space method): one WARP per problem.
//
// Replaces, per problem: QPNullSpaceSolver::Solve (source/qp.cc:679-729), which the NLS driver selects whenever a
// problem has equality constraints and no inequalities (source/nonlinear.cc:83-89, :248-258):
//     [Q1 Q2] R = A_eq^T P   (column-pivoted Householder QR),   u = Q1 R1^-T P^T (-b_eq),
//     (Q2^T G Q2) y = -Q2^T (c + G u)   (Cholesky of the reduced Hessian),   x = u + Q2 y.
// The pivoting rule, the Householder convention (beta = -sign(c0) |x|, tau = (beta - c0) / beta), the rank test and the
// Cholesky failure test (pivot <= 0 -> NOT_POSITIVE_DEFINITE) are those of Eigen 3.4.0's ColPivHouseholderQR / LLT, which
// is what the reference calls; the arithmetic is organised for a warp: lane = column for the reflector applications,
// lane = output element for the products, warp-shuffle reductions for the norms.
// Workspace (shared memory, or global scratch for the no-allocation entry point): QR (n x k), Q (n x n), G Z (n x nz),
// Z^T G Z (nz x nz), nz = n - k, and O(n) vectors: ns_smem_doubles(n, k).
#pragma once
#include <float.h>

#include "qp_common.cuh"

namespace mo {

struct NullSpaceArgs {
  int n, k;
  long long batch;
  const double* __restrict__ G;
  const double* __restrict__ c;
  const double* __restrict__ A;
  const double* __restrict__ b;
  double* __restrict__ x;
  int* __restrict__ status;
};

__host__ __device__ inline size_t ns_smem_doubles(int n, int k) {
  const int nz = n - k;
  return (size_t)n * k + (size_t)n * n + (size_t)n * nz + (size_t)nz * nz + 4 * (size_t)k + 3 * (size_t)n + nz + (k + 1) / 2 + 8;
}

__device__ __forceinline__ double ns_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Solves one problem held by pointers (global or shared). Returns the status to every lane; x (n) is written.
__device__ inline int ns_solve_warp(int n, int k, const double* __restrict__ G, const double* __restrict__ c,
                                    const double* __restrict__ A, const double* __restrict__ b, double* __restrict__ x,
                                    double* __restrict__ work) {
  const int lane = threadIdx.x & 31, rows = n, cols = k, size = n < k ? n : k, nz = n - k;
  double* qr = work;                       // rows x cols
  double* Q = qr + (size_t)rows * cols;    // rows x rows
  double* GZ = Q + (size_t)rows * rows;    // n x nz
  double* Gr = GZ + (size_t)n * nz;        // nz x nz
  double* hc = Gr + (size_t)nz * nz;       // k
  double* upd = hc + k;                    // k
  double* dir = upd + k;                   // k
  double* w = dir + k;                     // k
  double* u = w + k;                       // n
  double* gu = u + n;                      // n
  double* tmp = gu + n;                    // n
  double* y = tmp + n;                     // nz
  int* perm = reinterpret_cast<int*>(y + nz);  // k
#define NQR(i, j) qr[(size_t)(j) * rows + (i)]
#define NQM(i, j) Q[(size_t)(j) * rows + (i)]
  auto gsym = [&](int i, int j) { return i >= j ? G[(size_t)j * n + i] : G[(size_t)i * n + j]; };
  for (int e = lane; e < rows * cols; e += 32) {
    const int j = e / rows, i = e - j * rows;
    NQR(i, j) = A[(size_t)i * k + j];  // A^T
  }
  for (int e = lane; e < rows * rows; e += 32) Q[e] = (e / rows == e % rows) ? 1.0 : 0.0;
  __syncwarp();
  for (int j = lane; j < cols; j += 32) {
    double s = 0.0;
    for (int i = 0; i < rows; ++i) s += NQR(i, j) * NQR(i, j);
    upd[j] = dir[j] = sqrt(s);
    perm[j] = j;
  }
  __syncwarp();
  double maxnorm = 0.0;
  for (int j = 0; j < cols; ++j) maxnorm = fmax(maxnorm, upd[j]);
  const double th = maxnorm * DBL_EPSILON / (double)rows, threshold_helper = th * th, downdate = sqrt(DBL_EPSILON);
  int nonzero_pivots = size;
  double maxpivot = 0.0;
  for (int kk = 0; kk < size; ++kk) {
    int big = kk;
    double best = upd[kk];
    for (int j = kk + 1; j < cols; ++j)
      if (upd[j] > best) { best = upd[j]; big = j; }  // first maximum wins (uniform: every lane scans)
    if (nonzero_pivots == size && best * best < threshold_helper * (double)(rows - kk)) nonzero_pivots = kk;
    __syncwarp();
    if (big != kk) {
      for (int i = lane; i < rows; i += 32) { const double t = NQR(i, kk); NQR(i, kk) = NQR(i, big); NQR(i, big) = t; }
      if (lane == 0) {
        double t = upd[kk]; upd[kk] = upd[big]; upd[big] = t;
        t = dir[kk]; dir[kk] = dir[big]; dir[big] = t;
        const int pi = perm[kk]; perm[kk] = perm[big]; perm[big] = pi;
      }
    }
    __syncwarp();
    // makeHouseholderInPlace on column kk, rows kk.. (the tail norm is summed sequentially like Eigen's squaredNorm of a
    // short vector would not be: a tree sum here, the difference is rounding level)
    double tail = 0.0;
    for (int i = kk + 1 + lane; i < rows; i += 32) tail += NQR(i, kk) * NQR(i, kk);
    tail = ns_sum(tail);
    const double c0 = NQR(kk, kk);
    double beta, tau;
    __syncwarp();
    if (tail <= DBL_MIN) {
      tau = 0.0; beta = c0;
      for (int i = kk + 1 + lane; i < rows; i += 32) NQR(i, kk) = 0.0;
    } else {
      beta = sqrt(c0 * c0 + tail);
      if (c0 >= 0.0) beta = -beta;
      for (int i = kk + 1 + lane; i < rows; i += 32) NQR(i, kk) /= (c0 - beta);
      tau = (beta - c0) / beta;
    }
    if (lane == 0) { hc[kk] = tau; NQR(kk, kk) = beta; }
    maxpivot = fmax(maxpivot, fabs(beta));
    __syncwarp();
    // apply H = I - tau v v^T to the trailing columns (lane = column), then the norm downdate
    for (int j = kk + 1 + lane; j < cols; j += 32) {
      double t = 0.0;
      for (int i = kk + 1; i < rows; ++i) t += NQR(i, kk) * NQR(i, j);
      t += NQR(kk, j);
      NQR(kk, j) -= tau * t;
      for (int i = kk + 1; i < rows; ++i) NQR(i, j) -= (tau * NQR(i, kk)) * t;
      if (upd[j] != 0.0) {
        double r = fabs(NQR(kk, j)) / upd[j];
        r = (1.0 + r) * (1.0 - r);
        r = r < 0.0 ? 0.0 : r;
        const double ratio = upd[j] / dir[j], t2 = r * ratio * ratio;
        if (t2 <= downdate) {
          double s = 0.0;
          for (int i = kk + 1; i < rows; ++i) s += NQR(i, j) * NQR(i, j);
          dir[j] = sqrt(s);
          upd[j] = dir[j];
        } else {
          upd[j] *= sqrt(r);
        }
      }
    }
    __syncwarp();
  }
  int rank = 0;
  {
    const double pre = fabs(maxpivot) * (DBL_EPSILON * (double)size);
    for (int i = 0; i < nonzero_pivots; ++i) rank += fabs(NQR(i, i)) > pre;
  }
  for (int i = lane; i < n; i += 32) x[i] = 0.0;
  if (rank < k) { __syncwarp(); return 2; }
  // Q = H_0 ... H_{size-1} applied to the identity (lane = column of Q)
  for (int kk = size - 1; kk >= 0; --kk) {
    const double tau = hc[kk];
    for (int j = kk + lane; j < rows; j += 32) {
      double t = NQM(kk, j);
      for (int i = kk + 1; i < rows; ++i) t += NQR(i, kk) * NQM(i, j);
      NQM(kk, j) -= tau * t;
      for (int i = kk + 1; i < rows; ++i) NQM(i, j) -= (tau * NQR(i, kk)) * t;
    }
    __syncwarp();
  }
  // u = Q1 R1^-T P^T (-b): forward substitution with R1^T (sequential, lane 0), then the product (lane = row)
  if (lane == 0) {
    for (int i = 0; i < k; ++i) w[i] = -b[perm[i]];
    for (int i = 0; i < k; ++i) {
      double acc = w[i];
      for (int j = 0; j < i; ++j) acc -= NQR(j, i) * w[j];
      w[i] = acc / NQR(i, i);
    }
  }
  __syncwarp();
  for (int i = lane; i < n; i += 32) {
    double acc = 0.0;
    for (int j = 0; j < k; ++j) acc += NQM(i, j) * w[j];
    u[i] = acc;
  }
  __syncwarp();
  if (nz == 0) {
    for (int i = lane; i < n; i += 32) x[i] = u[i];
    __syncwarp();
    return 0;
  }
  for (int e = lane; e < n * nz; e += 32) {  // G Z
    const int j = e / n, i = e - j * n;
    double acc = 0.0;
    for (int l = 0; l < n; ++l) acc += gsym(i, l) * NQM(l, k + j);
    GZ[e] = acc;
  }
  __syncwarp();
  for (int e = lane; e < nz * nz; e += 32) {  // Z^T G Z
    const int j = e / nz, i = e - j * nz;
    double acc = 0.0;
    for (int l = 0; l < n; ++l) acc += NQM(l, k + i) * GZ[(size_t)j * n + l];
    Gr[e] = acc;
  }
  for (int i = lane; i < n; i += 32) {  // c + G u
    double acc = c[i];
    for (int l = 0; l < n; ++l) acc += gsym(i, l) * u[l];
    gu[i] = acc;
  }
  __syncwarp();
  // LLT<Lower> (left-looking by column, lane = row below the diagonal)
  for (int kk = 0; kk < nz; ++kk) {
    double d = Gr[(size_t)kk * nz + kk];
    for (int j = 0; j < kk; ++j) d -= Gr[(size_t)j * nz + kk] * Gr[(size_t)j * nz + kk];
    if (d <= 0.0) { __syncwarp(); return 1; }  // uniform: every lane computes the same d
    d = sqrt(d);
    __syncwarp();
    if (lane == 0) Gr[(size_t)kk * nz + kk] = d;
    for (int i = kk + 1 + lane; i < nz; i += 32) {
      double acc = Gr[(size_t)kk * nz + i];
      for (int j = 0; j < kk; ++j) acc -= Gr[(size_t)j * nz + i] * Gr[(size_t)j * nz + kk];
      Gr[(size_t)kk * nz + i] = acc / d;
    }
    __syncwarp();
  }
  for (int j = lane; j < nz; j += 32) {
    double acc = 0.0;
    for (int l = 0; l < n; ++l) acc += NQM(l, k + j) * gu[l];
    y[j] = -acc;
  }
  __syncwarp();
  if (lane == 0) {
    for (int i = 0; i < nz; ++i) {
      double acc = y[i];
      for (int j = 0; j < i; ++j) acc -= Gr[(size_t)j * nz + i] * y[j];
      y[i] = acc / Gr[(size_t)i * nz + i];
    }
    for (int i = nz - 1; i >= 0; --i) {
      double acc = y[i];
      for (int j = i + 1; j < nz; ++j) acc -= Gr[(size_t)i * nz + j] * y[j];
      y[i] = acc / Gr[(size_t)i * nz + i];
    }
  }
  __syncwarp();
  for (int i = lane; i < n; i += 32) {
    double acc = u[i];
    for (int j = 0; j < nz; ++j) acc += NQM(i, k + j) * y[j];
    x[i] = acc;
  }
  __syncwarp();
#undef NQR
#undef NQM
  return 0;
}

static __global__ void __launch_bounds__(32) qp_nullspace_kernel(const NullSpaceArgs a) {
  extern __shared__ double ns_smem[];
  const int n = a.n, k = a.k;
  for (long long p = blockIdx.x; p < a.batch; p += gridDim.x) {
    __syncwarp();
    const int st = ns_solve_warp(n, k, a.G + p * (long long)n * n, a.c + p * n, a.A + p * (long long)n * k, a.b + p * k,
                                 a.x + p * n, ns_smem);
    if (threadIdx.x == 0) a.status[p] = st;
  }
}

}  // namespace mo
