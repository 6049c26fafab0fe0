// Batched nonlinear least squares on the device (sm_100a), one WARP per problem:
//   nls_linearize_kernel     LinearizeAndFillQP (source/nonlinear.cc:170-214) + Residual::UpdateHessian / UpdateJacobian
//                            (include/mini_opt/residual.hpp:186-250) for dense per-problem Jacobians
//   chain_fk_kernel          ActuatorChain::Update (test/transform_chains.cc:165-244)
//   nls_chain_ik_kernel      ConstrainedNonlinearLeastSquares::Solve (source/nonlinear.cc:75-158) for the transform-chain
//                            residual family: FK + analytic Jacobians, J'J / J'r + lambda I, the interior point QP
//                            (generic 32-thread solver, QP held in shared memory), directional derivatives, L1 merit
//                            penalty, Armijo / polynomial line search, LM lambda state machine.
// The forward kinematics are organised for a warp: lane i evaluates link i (the trigonometry), lanes 0 and 1 run the
// two sequential scans (suffix poses, prefix rotations), lane i then forms link i's Jacobian columns.
#pragma once

#include "nls_launch.h"
#include "qp_generic.cuh"
#include "qp_warp.cuh"
#include "qp_nullspace.cuh"

namespace mo {


struct ChainArgs {
  int links, n, n_cost, cost_block, n_eq, m;
  long long batch;
  const double* __restrict__ link_pose;
  const unsigned char* __restrict__ link_mask;
  const double* __restrict__ cost_w;
  const double* __restrict__ cost_target;
  const double* __restrict__ eq_w;
  const double* __restrict__ eq_target;
  const int* __restrict__ ineq_var;
  const double* __restrict__ ineq_a;
  const double* __restrict__ ineq_b;
  const double* __restrict__ x0;
};

struct NlsParams {  // mo_nls_params
  int max_iterations, max_qp_iterations;
  double kkt_tol, absolute_exit_tol, relative_exit_tol, absolute_first_derivative_tol;
  int max_line_search_iterations, line_search_strategy;
  double armijo_search_tau, equality_penalty_initial, equality_penalty_scale_factor, equality_penalty_rho;
  double lambda_initial, lambda_failure_init, lambda_decrease_on_success, lambda_decrease_on_restore, max_lambda, min_lambda;
  int retraction, inner_qp_solver;
};

struct NlsOut {
  double* __restrict__ x;
  int* __restrict__ termination;
  int* __restrict__ iterations;
  int* __restrict__ qp_iterations;
  int* __restrict__ ls_steps;
  double* __restrict__ errors;
};

struct Quat { double w, x, y, z; };

__device__ __forceinline__ Quat qmul(const Quat& a, const Quat& b) {
  Quat r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  return r;
}

// column-major 3x3 rotation matrix of a unit quaternion
__device__ __forceinline__ void qmatrix(const Quat& q, double* R) {
  const double xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
  const double xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
  const double wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
  R[0] = 1 - 2 * (yy + zz); R[3] = 2 * (xy - wz);     R[6] = 2 * (xz + wy);
  R[1] = 2 * (xy + wz);     R[4] = 1 - 2 * (xx + zz); R[7] = 2 * (yz - wx);
  R[2] = 2 * (xz - wy);     R[5] = 2 * (yz + wx);     R[8] = 1 - 2 * (xx + yy);
}

// Shared-memory state of one chain (one warp).
struct ChainWork {
  int links, n;
  double* pose;    // [links][7] copy of link_pose
  double* base;    // [links][3] rotation_xyz of the ActuatorLink constructor (transform_chains.cc:107-122)
  double* hsc;     // [links][6] (sin, cos) of base / 2 per axis: the half-angle factors of the axes that are NOT parameters
  double* lq;      // [links][4] link rotations at the current parameters
  double* lt;      // [links][3] link translations
  double* lD;      // [links][9] d(right tangent)/d(euler xyz)
  double* sufq;    // [links+1][4]  i_R_end
  double* suft;    // [links+1][3]  i_t_end
  double* preR;    // [links+1][9]  start_R_i as matrices
  double* rotD;    // [3n] rotation_D_params (column-major 3 x n)
  double* transD;  // [3n] translation_D_params
  double* cached;  // [n]  params of the last update (ShouldUpdate, transform_chains.cc:248-258)
  int* pos;        // [links] first parameter index of each link
  unsigned char* mask;  // [links][6]
  int* have_cache;
};

__host__ __device__ inline size_t chain_smem_doubles(int links, int n) {
  size_t d = (size_t)links * (7 + 3 + 6 + 4 + 3 + 9) + (size_t)(links + 1) * (4 + 3 + 9) + 7 * (size_t)n;
  d += (size_t)links / 2 + 1;           // pos
  d += ((size_t)links * 6 + 7) / 8 + 1;  // mask
  d += 1;                               // have_cache
  return d;
}

__device__ inline double* chain_carve(ChainWork& C, double* q, int links, int n) {
  C.links = links; C.n = n;
  C.pose = q; q += links * 7;
  C.base = q; q += links * 3;
  C.hsc = q; q += links * 6;
  C.lq = q; q += links * 4;
  C.lt = q; q += links * 3;
  C.lD = q; q += links * 9;
  C.sufq = q; q += (links + 1) * 4;
  C.suft = q; q += (links + 1) * 3;
  C.preR = q; q += (links + 1) * 9;
  C.rotD = q; q += 3 * n;
  C.transD = q; q += 3 * n;
  C.cached = q; q += n;
  C.pos = reinterpret_cast<int*>(q); q += links / 2 + 1;
  C.mask = reinterpret_cast<unsigned char*>(q); q += (links * 6 + 7) / 8 + 1;
  C.have_cache = reinterpret_cast<int*>(q); q += 1;
  return q;
}

// Chain description into shared memory; Euler decomposition of the links with an active rotation.
__device__ inline void chain_setup(ChainWork& C, const ChainArgs& a) {
  const int lane = threadIdx.x & 31;
  for (int e = lane; e < C.links * 7; e += 32) C.pose[e] = a.link_pose[e];
  for (int e = lane; e < C.links * 6; e += 32) C.mask[e] = a.link_mask[e];
  __syncwarp();
  if (lane == 0) {
    int p = 0;
    for (int i = 0; i < C.links; ++i) {
      C.pos[i] = p;
      for (int j = 0; j < 6; ++j) p += C.mask[6 * i + j] > 0;
    }
    *C.have_cache = 0;
  }
  for (int i = lane; i < C.links; i += 32) {
    const double* P = C.pose + 7 * i;
    // -EulerAnglesFromSO3(q.conjugate()), so3.hpp:365-371
    const double w = P[0], x = -P[1], y = -P[2], z = -P[3];
    C.base[3 * i] = -atan2(2 * (y * z + x * w), 1 - 2 * (x * x + y * y));
    C.base[3 * i + 1] = -asin(fmax(fmin(1., -2 * (x * z - y * w)), -1.));
    C.base[3 * i + 2] = -atan2(2 * (x * y + z * w), 1 - 2 * (y * y + z * z));
    // constant axes: the same sincos(angle / 2) chain_update would evaluate at every call (bit-identical, evaluated once)
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) sincos(C.base[3 * i + ax] / 2, &C.hsc[6 * i + 2 * ax], &C.hsc[6 * i + 2 * ax + 1]);
  }
  for (int e = lane; e < 3 * C.n; e += 32) { C.rotD[e] = 0.0; C.transD[e] = 0.0; }
  __syncwarp();
}

// ActuatorChain::Update (transform_chains.cc:165-244) with ComputeChain (:23-82), executed by one warp.
// Returns without work when no parameter moved by more than 1e-9 since the last update (the reference's cache).
__device__ inline void chain_update(ChainWork& C, const double* params) {
  const int lane = threadIdx.x & 31, L = C.links, n = C.n;
  bool changed = false;
  for (int i = lane; i < n; i += 32) changed = changed || fabs(C.cached[i] - params[i]) > 1.0e-9;
  const bool had = *C.have_cache != 0;
  if (had && !__any_sync(0xffffffffu, changed)) return;
  __syncwarp();
  for (int i = lane; i < n; i += 32) C.cached[i] = params[i];
  if (lane == 0) *C.have_cache = 1;
  // (1) link poses at the current parameters: ActuatorLink::Compute (:125-158), SO3FromEulerAngles XYZ (so3.hpp:318-333)
  for (int i = lane; i < L; i += 32) {
    const unsigned char* act = C.mask + 6 * i;
    const double* P = C.pose + 7 * i;
    double pu[6] = {C.base[3 * i], C.base[3 * i + 1], C.base[3 * i + 2], P[4], P[5], P[6]};
    int pi = C.pos[i];
    const bool rot = (act[0] | act[1] | act[2]) != 0;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (act[k] && (rot || k >= 3)) pu[k] = params[pi++];
    Quat q;
    double* D = C.lD + 9 * i;
    if (rot) {
      // an axis that is a parameter moves; the others keep the half-angle factors computed in chain_setup
      const double* H = C.hsc + 6 * i;
      double s_hx = H[0], c_hx = H[1], s_hy = H[2], c_hy = H[3], s_hz = H[4], c_hz = H[5];
      if (act[0]) sincos(pu[0] / 2, &s_hx, &c_hx);
      if (act[1]) sincos(pu[1] / 2, &s_hy, &c_hy);
      if (act[2]) sincos(pu[2] / 2, &s_hz, &c_hz);
      const double cy = 2 * c_hy * c_hy - 1, sy = 2 * c_hy * s_hy;
      const double cz = 2 * c_hz * c_hz - 1, sz = 2 * c_hz * s_hz;
      q.w = c_hx * c_hy * c_hz - s_hx * s_hy * s_hz;
      q.x = c_hx * s_hy * s_hz + c_hy * c_hz * s_hx;
      q.y = c_hx * c_hz * s_hy - c_hy * s_hx * s_hz;
      q.z = c_hx * c_hy * s_hz + c_hz * s_hx * s_hy;
      D[0] = cy * cz; D[1] = -sz * cy; D[2] = sy;
      D[3] = sz; D[4] = cz; D[5] = 0.0;
      D[6] = 0.0; D[7] = 0.0; D[8] = 1.0;
    } else {
      q.w = P[0]; q.x = P[1]; q.y = P[2]; q.z = P[3];
    }
    C.lq[4 * i] = q.w; C.lq[4 * i + 1] = q.x; C.lq[4 * i + 2] = q.y; C.lq[4 * i + 3] = q.z;
    C.lt[3 * i] = pu[3]; C.lt[3 * i + 1] = pu[4]; C.lt[3 * i + 2] = pu[5];
  }
  __syncwarp();
  // (2) the two scans: lane 0 suffix poses i_R_end / i_t_end (:36-52), lane 1 prefix rotations start_R_i (:56-61)
  if (lane == 0) {
    Quat s{1, 0, 0, 0};
    double t[3] = {0, 0, 0};
    C.sufq[4 * L] = 1; C.sufq[4 * L + 1] = 0; C.sufq[4 * L + 2] = 0; C.sufq[4 * L + 3] = 0;
    C.suft[3 * L] = 0; C.suft[3 * L + 1] = 0; C.suft[3 * L + 2] = 0;
    for (int i = L - 1; i >= 0; --i) {
      const Quat q{C.lq[4 * i], C.lq[4 * i + 1], C.lq[4 * i + 2], C.lq[4 * i + 3]};
      double R[9];
      qmatrix(q, R);
      const double t0 = R[0] * t[0] + R[3] * t[1] + R[6] * t[2] + C.lt[3 * i];
      const double t1 = R[1] * t[0] + R[4] * t[1] + R[7] * t[2] + C.lt[3 * i + 1];
      const double t2 = R[2] * t[0] + R[5] * t[1] + R[8] * t[2] + C.lt[3 * i + 2];
      t[0] = t0; t[1] = t1; t[2] = t2;
      s = qmul(q, s);
      C.sufq[4 * i] = s.w; C.sufq[4 * i + 1] = s.x; C.sufq[4 * i + 2] = s.y; C.sufq[4 * i + 3] = s.z;
      C.suft[3 * i] = t0; C.suft[3 * i + 1] = t1; C.suft[3 * i + 2] = t2;
    }
  } else if (lane == 1) {
    Quat p{1, 0, 0, 0};
    for (int i = 0; i <= L; ++i) {
      qmatrix(p, C.preR + 9 * i);
      if (i < L) p = qmul(p, Quat{C.lq[4 * i], C.lq[4 * i + 1], C.lq[4 * i + 2], C.lq[4 * i + 3]});
    }
  }
  __syncwarp();
  // (3) Jacobian columns of link i (:66-81 and the chain rule :198-243)
  for (int i = lane; i < L; i += 32) {
    const unsigned char* act = C.mask + 6 * i;
    const int nr = (act[0] > 0) + (act[1] > 0) + (act[2] > 0);
    const int ac = nr + (act[3] > 0) + (act[4] > 0) + (act[5] > 0);
    if (ac == 0) continue;
    const int pos = C.pos[i];
    if (nr > 0) {
      double tDr[9], rDr[9];
      if (i < L - 1) {
        const double* Pn = C.preR + 9 * (i + 1);
        const double v0 = -C.suft[3 * (i + 1)], v1 = -C.suft[3 * (i + 1) + 1], v2 = -C.suft[3 * (i + 1) + 2];
        // P * Skew3(v): column c of Skew3(v) is e_c x ... : S = [[0,-v2,v1],[v2,0,-v0],[-v1,v0,0]]
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const double p0 = Pn[r], p1 = Pn[3 + r], p2 = Pn[6 + r];
          tDr[r] = p1 * v2 - p2 * v1;
          tDr[3 + r] = p2 * v0 - p0 * v2;
          tDr[6 + r] = p0 * v1 - p1 * v0;
        }
        const Quat sc{C.sufq[4 * (i + 1)], -C.sufq[4 * (i + 1) + 1], -C.sufq[4 * (i + 1) + 2], -C.sufq[4 * (i + 1) + 3]};
        qmatrix(sc, rDr);
      } else {
#pragma unroll
        for (int k = 0; k < 9; ++k) { tDr[k] = 0.0; rDr[k] = (k % 4 == 0) ? 1.0 : 0.0; }
      }
      const double* D = C.lD + 9 * i;
      int oi = 0;
#pragma unroll
      for (int axis = 0; axis < 3; ++axis) {
        if (!act[axis]) continue;
        const double d0 = D[3 * axis], d1 = D[3 * axis + 1], d2 = D[3 * axis + 2];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          C.transD[3 * (pos + oi) + r] = tDr[r] * d0 + tDr[3 + r] * d1 + tDr[6 + r] * d2;
          C.rotD[3 * (pos + oi) + r] = rDr[r] * d0 + rDr[3 + r] * d1 + rDr[6 + r] * d2;
        }
        ++oi;
      }
    }
    if (ac - nr > 0) {
      const double* Pi = C.preR + 9 * i;
      int oi = 0;
#pragma unroll
      for (int axis = 0; axis < 3; ++axis) {
        if (!act[axis + 3]) continue;
#pragma unroll
        for (int r = 0; r < 3; ++r) C.transD[3 * (pos + nr + oi) + r] = Pi[3 * axis + r];
        ++oi;
      }
    }
  }
  __syncwarp();
}

// mo_chain_fk_batch: one warp per parameter vector.
static __global__ void __launch_bounds__(32) chain_fk_kernel(const ChainArgs a, const double* __restrict__ params,
                                                      double* __restrict__ translation, double* __restrict__ rotation_q,
                                                      double* __restrict__ tD, double* __restrict__ rD) {
  extern __shared__ double smem[];
  ChainWork C;
  chain_carve(C, smem, a.links, a.n);
  chain_setup(C, a);
  const int lane = threadIdx.x & 31, n = a.n;
  for (long long p = blockIdx.x; p < a.batch; p += gridDim.x) {
    if (lane == 0) *C.have_cache = 0;
    __syncwarp();
    chain_update(C, params + p * n);
    if (translation && lane < 3) translation[3 * p + lane] = C.suft[lane];
    if (rotation_q && lane < 4) rotation_q[4 * p + lane] = C.sufq[lane];
    for (int e = lane; e < 3 * n; e += 32) {
      if (tD) tD[3 * n * p + e] = C.transD[e];
      if (rD) rD[3 * n * p + e] = C.rotD[e];
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------
// Gauss-Newton / LM linearization (LinearizeAndFillQP, nonlinear.cc:170-214), executed by one warp.
// Jc: r_cost x n column-major (leading dimension r_cost), Je: k x n column-major. G n x n column-major: lower triangle
// = sum over cost blocks of J_b' J_b (+ lambda on the diagonal), strict upper triangle = 0 like the reference.
// errors = {f, equality}. All output pointers may be shared or global memory.
// ---------------------------------------------------------------------------------------------------------
__device__ inline void linearize_warp(int n, int r_cost, int cost_block, int k, int m, const double* x, const double* Jc,
                                      const double* rc, const double* Je, const double* re, double lambda,
                                      const int* ivar, const double* ia, const double* ib, double* G, double* c,
                                      double* A, double* b, double* ib_shift, double* errors) {
  const int lane = threadIdx.x & 31;
  const int cb = cost_block > 0 ? cost_block : r_cost;
  for (int e = lane; e < n * n; e += 32) {
    const int col = e / n, row = e - col * n;
    double g = 0.0;
    if (row >= col) {
      for (int r0 = 0; r0 < r_cost; r0 += cb) {  // one Residual block at a time (UpdateHessian, residual.hpp:199-224)
        double d = 0.0;
        for (int i = 0; i < cb; ++i) d += Jc[row * r_cost + r0 + i] * Jc[col * r_cost + r0 + i];
        g += d;
      }
      if (row == col && lambda > 0) g += lambda;
    }
    G[e] = g;
  }
  for (int row = lane; row < n; row += 32) {
    double g = 0.0;
    for (int r0 = 0; r0 < r_cost; r0 += cb) {
      double d = 0.0;
      for (int i = 0; i < cb; ++i) d += Jc[row * r_cost + r0 + i] * rc[r0 + i];
      g += d;
    }
    c[row] = g;
  }
  for (int e = lane; e < k * n; e += 32) A[e] = Je[e];
  for (int r = lane; r < k; r += 32) b[r] = re[r];
  for (int i = lane; i < m; i += 32) ib_shift[i] = ia[i] * x[ivar[i]] + ib[i];  // ShiftTo, qp.hpp:57-59
  if (lane == 0) {
    double f = 0.0, eq = 0.0;
    for (int r0 = 0; r0 < r_cost; r0 += cb) {
      double s = 0.0;
      for (int i = 0; i < cb; ++i) s += rc[r0 + i] * rc[r0 + i];
      f += 0.5 * s;
    }
    for (int r = 0; r < k; ++r) eq += fabs(re[r]);
    errors[0] = f;
    errors[1] = eq;
  }
  __syncwarp();
}

struct LinearizeArgs {
  int n, r_cost, cost_block, r_eq, m;
  long long batch;
  const double *x, *J_cost, *r_cost_v, *J_eq, *r_eq_v, *lambda;
  const int* ineq_var;
  const double *ineq_a, *ineq_b;
  double *G, *c, *A_eq, *b_eq, *ineq_b_shifted, *errors;
};

// mo_nls_linearize_batch: one warp per problem, 4 warps per block.
static __global__ void __launch_bounds__(32) nls_linearize_kernel(const LinearizeArgs a) {
  const int n = a.n, k = a.r_eq, m = a.m;
  for (long long p = blockIdx.x; p < a.batch; p += gridDim.x) {
    linearize_warp(n, a.r_cost, a.cost_block, k, m, a.x + p * n, a.J_cost + p * a.r_cost * n, a.r_cost_v + p * a.r_cost,
                   k ? a.J_eq + p * k * n : nullptr, k ? a.r_eq_v + p * k : nullptr, a.lambda[p], a.ineq_var + p * m,
                   a.ineq_a + p * m, a.ineq_b + p * m, a.G + p * n * n, a.c + p * n, k ? a.A_eq + p * k * n : nullptr,
                   k ? a.b_eq + p * k : nullptr, a.ineq_b_shifted + p * m, a.errors + 2 * p);
  }
}

// ---------------------------------------------------------------------------------------------------------
// The SQP + LM loop for the chain family
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double mod_pi(double angle) {  // angle_utils.hpp:9-17
  const double pi = 3.14159265358979323846, two_pi = 2 * pi;
  angle = fmod(angle, two_pi);
  angle += (angle < 0) ? two_pi : 0.0;
  angle -= (angle > pi) ? two_pi : 0.0;
  return angle;
}

// rows w_i . t - target_i (and J rows w_i' * translation_D_params) at parameters x
__device__ inline void chain_rows(ChainWork& C, const double* x, int rows, const double* w, const double* target,
                                  double* r, double* J) {
  chain_update(C, x);
  const int lane = threadIdx.x & 31, n = C.n;
  for (int i = lane; i < rows; i += 32)
    r[i] = (w[3 * i] * C.suft[0] + w[3 * i + 1] * C.suft[1]) + w[3 * i + 2] * C.suft[2] - target[i];
  if (J)
    for (int e = lane; e < rows * n; e += 32) {
      const int c = e / rows, i = e - c * rows;
      J[e] = (w[3 * i] * C.transD[3 * c] + w[3 * i + 1] * C.transD[3 * c + 1]) + w[3 * i + 2] * C.transD[3 * c + 2];
    }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------
// Residual FAMILY: what the SQP kernel needs from a problem class (the device-side counterpart of the reference's
// type-erased Residual, include/mini_opt/residual.hpp:25-143). A family is a struct with
//   Args   POD kernel argument with at least { n, n_cost, cost_block, n_eq, m, batch, ineq_var, ineq_a, ineq_b, x0 }
//   Work   per-warp working state (pointers into the warp's shared-memory region)
//   smem_doubles(args), carve(work, q, args) -> q', setup(work, args)   once per warp
//   begin(work, args, p)                                              a new problem starts in this warp
//   cost_rows / eq_rows(work, args, p, x, r, J)    residual rows at x (and the column-major rows x n Jacobian unless
//                                                  J == nullptr), written to shared memory, warp-synchronous
// ChainFamily is the transform-chain IK family; nls_families.cuh adds functor-defined families.
// ---------------------------------------------------------------------------------------------------------
struct ChainFamily {
  using Args = ChainArgs;
  using Work = ChainWork;
  __host__ __device__ static size_t smem_doubles(const Args& a) { return chain_smem_doubles(a.links, a.n); }
  __device__ static double* carve(Work& C, double* q, const Args& a) { return chain_carve(C, q, a.links, a.n); }
  __device__ static void setup(Work& C, const Args& a) { chain_setup(C, a); }
  __device__ static void begin(Work& C, const Args&, long long) {
    if ((threadIdx.x & 31) == 0) *C.have_cache = 0;
  }
  __device__ static void cost_rows(Work& C, const Args& a, long long p, const double* x, double* r, double* J) {
    chain_rows(C, x, a.n_cost, a.cost_w, a.cost_target + p * a.n_cost, r, J);
  }
  __device__ static void eq_rows(Work& C, const Args& a, long long p, const double* x, double* r, double* J) {
    chain_rows(C, x, a.n_eq, a.eq_w, a.eq_target + p * a.n_eq, r, J);
  }
};

// EvaluateNonlinearErrors (nonlinear.cc:279-293): every lane returns {f, equality}
template <class F>
__device__ inline void family_errors(typename F::Work& C, const typename F::Args& a, long long p, const double* x,
                                     double* rbuf, double& f, double& eq) {
  const int cb = a.cost_block > 0 ? a.cost_block : a.n_cost;
  F::cost_rows(C, a, p, x, rbuf, nullptr);
  f = 0.0;
  for (int r0 = 0; r0 < a.n_cost; r0 += cb) {
    double s = 0.0;
    for (int i = 0; i < cb; ++i) s += rbuf[r0 + i] * rbuf[r0 + i];
    f += 0.5 * s;
  }
  eq = 0.0;
  __syncwarp();
  if (a.n_eq) {
    F::eq_rows(C, a, p, x, rbuf, nullptr);
    for (int i = 0; i < a.n_eq; ++i) eq += fabs(rbuf[i]);
    __syncwarp();
  }
}

__host__ __device__ inline size_t nls_smem_doubles(size_t family_doubles, int n, int n_cost, int k, int m) {
  // inner QP workspace: the interior point solver's, or (equalities and no inequalities) the null-space solver's
  const size_t qpw = (m == 0 && k > 0 && ns_smem_doubles(n, k) > gen_smem_doubles(n, k, m, 32)) ? ns_smem_doubles(n, k)
                                                                                               : gen_smem_doubles(n, k, m, 32);
  size_t d = family_doubles + qpw;
  d += (size_t)n * n + n + (size_t)k * n + k + m;       // QP: G, c, A, b, shifted bounds
  d += (size_t)n_cost * n + (size_t)k * n + 3 * (size_t)n; // Jc, Je, x, cand, dx
  d += 2 * kMaxChainRows + 8;                            // residual buffers, errors
  return d;
}

// ConstrainedNonlinearLeastSquares::Solve (nonlinear.cc:75-158), one warp per problem. Resident warps per SM, measured on
// B200 in round 1 with one warp per CTA (100K problems, 6 links / 12 links, problems/s): 12 -> 700K / 229K, 16 -> 787K / 229K,
// 24 -> 802K / 203K (80 registers: spills at n = 12), 32 -> 794K / 179K: the launch bounds below keep W x MINB = 16.
// Problems are handed out through a global ticket counter (zeroed on the launch stream): SQP iteration counts are heavy-tailed
// (a few % of the IK problems run to max_iterations), a static stride leaves most warps idle while the unlucky ones finish.
constexpr int kIkTicketSlots = 64;  // launches in flight that may hold a ticket counter at the same time
static __device__ unsigned long long g_ik_tickets[kIkTicketSlots];  // one ring per translation unit

// LOCK-STEP schedule: W warps per CTA, each warp owns one problem at a time (ticket queue), and the warps of a CTA walk the
// phases of an SQP iteration TOGETHER: [fetch] | linearize (FK + Jacobians + J'J) | inner QP | derivatives + line search +
// lambda update, one CTA barrier after each. Why: the one-warp-per-CTA kernel walks ~5K instructions = 80 KB of text per SQP
// iteration against a 32 KB L1.5 instruction cache with its 16 resident warps each somewhere else in that text -- measured
// ~28 cycles per issued warp instruction, an L2 round trip per fetched line (profiles/r01_ik6_ncu_full_summary.txt:
// stall_no_instruction 9.8 per issue). In lock-step a fetched line serves W warps, and the loops inside a phase (Newton steps of
// the QP, line search evaluations) have bodies that fit the cache once nobody else is evicting them. The arithmetic of a
// problem is untouched: results are bit-identical to the W = 1 schedule. Cost: warps idle at the barriers while the slowest
// QP / line search of the CTA finishes.
#ifdef MO_IK_TRACE
// Debug build only: lane 0 of warp 0 of CTA 0 accumulates cycles per phase of the lock-step pass and prints them at kernel end.
__shared__ long long s_ik_tr[8];
__shared__ long long s_ik_t0;
#define IK_T(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long t_ = clock64(); s_ik_tr[i] += t_ - s_ik_t0; s_ik_t0 = t_; } } while (0)
#else
#define IK_T(i) do {} while (0)
#endif

template <class F, int W, int MINB>
__global__ void __launch_bounds__(32 * W, MINB) nls_solve_kernel(const typename F::Args a, const NlsParams prm, const NlsOut out,
                                                                 unsigned long long* __restrict__ ticket,
                                                                 const long long per_warp_doubles, const int sync_mask) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n = a.n, k = a.n_eq, m = a.m;
  typename F::Work C;
  double* q = F::carve(C, smem + (long long)warp * per_warp_doubles, a);
  GenWork W_;
  gen_carve(W_, q, n, k, m, 32);
  const bool null_space = m == 0 && k > 0;  // solver choice of the reference, nonlinear.cc:83-89
  double* nswork = q;                       // QPNullSpaceSolver workspace aliases the interior point workspace
  q += (null_space && ns_smem_doubles(n, k) > gen_smem_doubles(n, k, m, 32)) ? ns_smem_doubles(n, k) : gen_smem_doubles(n, k, m, 32);
  double* G = q; q += n * n;
  double* cvec = q; q += n;
  double* A = q; q += k * n;
  double* bvec = q; q += k;
  double* Jc = q; q += a.n_cost * n;
  double* Je = q; q += k * n;
  double* x = q; q += n;
  double* cand = q; q += n;
  double* dx = q; q += n;
  double* rc = q; q += kMaxChainRows;
  double* re = q; q += kMaxChainRows;
  double* err = q; q += 8;
  F::setup(C, a);
  for (int i = lane; i < m; i += 32) { W_.cv[i] = a.ineq_var[i]; W_.ca[i] = a.ineq_a[i]; }
  __syncwarp();
  GenProb Q{G, cvec, k ? A : nullptr, k ? bvec : nullptr};
  const GenSolveParams sp{1.0, 0.1, prm.kkt_tol, 1.0e-6, prm.max_qp_iterations, MO_BARRIER_COMPLEMENTARITY, 0,
                          k ? MO_GUESS_SOLVE_EQUALITY_CONSTRAINED : MO_GUESS_NAIVE, 0};  // nonlinear.cc:226-239
  constexpr int kMaxSteps = 40;
  // state of this warp's problem (uniform over the lanes)
  long long p = -1;
  bool busy = false, drained = false;
  int pass_no = 0;
  int state = 0, term = MO_NLS_MAX_ITERATIONS, iters = 0, qp_total = 0, ls_total = 0, iter = 0;
  double lambda = 0.0, penalty = 0.0;
  // results of problem p: x, the errors at x, the counters
  auto finish = [&]() {
    __syncwarp();
    for (int i = lane; i < n; i += 32) out.x[p * n + i] = x[i];
    double ff = 0.0, fe = 0.0;
    if (out.errors) family_errors<F>(C, a, p, x, rc, ff, fe);
    if (lane == 0) {
      out.termination[p] = term;
      out.iterations[p] = iters;
      if (out.qp_iterations) out.qp_iterations[p] = qp_total;
      if (out.ls_steps) out.ls_steps[p] = ls_total;
      if (out.errors) { out.errors[2 * p] = ff; out.errors[2 * p + 1] = fe; }
    }
    __syncwarp();
    busy = false;
  };
#ifdef MO_IK_TRACE
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) s_ik_tr[i] = 0; s_ik_t0 = clock64(); }
#endif
  for (;;) {
    IK_T(6);  // finish() of the previous pass
    // ---- fetch: an idle warp takes the next problem of the batch ----
    while (!busy && !drained) {
      unsigned long long t = 0;
      if (lane == 0) t = atomicAdd(ticket, 1ULL);
      p = (long long)__shfl_sync(0xffffffffu, t, 0);
      if (p >= a.batch) { drained = true; break; }
      F::begin(C, a, p);
      for (int i = lane; i < n; i += 32) x[i] = a.x0[p * n + i];
      __syncwarp();
      state = 0; term = MO_NLS_MAX_ITERATIONS; iters = 0; qp_total = 0; ls_total = 0; iter = 0;
      lambda = prm.lambda_initial; penalty = prm.equality_penalty_initial;
      busy = true;
      if (prm.max_iterations <= 0) finish();  // the Solve loop never runs: MAX_ITERATIONS at x0
    }
    IK_T(0);  // fetch
    // sync_mask bits 2-3: the opening barrier (which is also the CTA's "anyone still busy" vote) only every 2nd / 4th pass;
    // in between the warps run on, a drained warp idles until the next vote
    if (W > 1) {
      const int skip = (sync_mask >> 2) & 3;
      if ((pass_no++ & skip) == 0) { if (!__syncthreads_or(busy ? 1 : 0)) break; }
    } else if (!busy) break;
    IK_T(1);  // wait at the barrier that opens the pass
    bool ended = false;  // this problem terminates in this pass
    double pre_f = 0.0, pre_eq = 0.0;
    // ---- phase A: LinearizeAndFillQP (nonlinear.cc:170-214) ----
    if (busy) {
      F::cost_rows(C, a, p, x, rc, Jc);
      if (k) F::eq_rows(C, a, p, x, re, Je);
      linearize_warp(n, a.n_cost, a.cost_block, k, m, x, Jc, rc, Je, re, lambda, a.ineq_var, a.ineq_a, a.ineq_b, G, cvec,
                     A, bvec, W_.cb, err);
      pre_f = err[0]; pre_eq = err[1];
    }
    IK_T(2);  // phase A
    if (W > 1 && (sync_mask & 1)) __syncthreads();  // uniform over the CTA: a kernel argument
    IK_T(3);  // wait after A
    // ---- phase B: ComputeStepDirection (nonlinear.cc:216-259): Setup + Solve on the QP in shared memory ----
    if (busy) {
      for (int i = lane; i < W_.T; i += 32) W_.vars[i] = 0.0;
      __syncwarp();
      int qiters = 0;
      double kkt_after[4];
      if (null_space) {  // QPNullSpaceSolver (qp.cc:679-729); dx = 0 unless SUCCESS (nonlinear.cc:248-258)
        const int ns = ns_solve_warp(n, k, G, cvec, A, bvec, dx, nswork);
        if (ns == 1) { term = MO_NLS_QP_INDEFINITE; ended = true; }  // QPWasIndefinite, nonlinear.cc:103-105
        else if (ns != 0) { term = MO_NLS_QP_FAILED; ended = true; }
        else qp_total += 1;  // NumQPIterations counts one per null-space solve (structs.cc:208-219)
      } else {
        int qstatus = prm.inner_qp_solver == MO_NLS_QP_GENERIC ? -1 : wip_solve_any(W_, Q, sp, qiters, kkt_after);
        if (qstatus < 0) qstatus = gen_solve<32>(W_, Q, sp, nullptr, qiters, kkt_after);
        if (qstatus >= MO_QP_FAILED_FACTORIZATION) { term = MO_NLS_QP_FAILED; ended = true; }  // the reference throws here
        else {
          qp_total += qiters;
          for (int i = lane; i < n; i += 32) dx[i] = W_.vars[i];
        }
      }
      __syncwarp();
    }
    IK_T(4);  // phase B
    if (W > 1 && (sync_mask & 2)) __syncthreads();
    // ---- phase C: derivatives, penalty, line search, lambda update ----
    if (busy && !ended) {
      // ComputeQPCostDerivative (:457-482), SelectPenalty (:484-502)
      double d_f = 0.0, d_eq = 0.0;
      for (int i = 0; i < n; ++i) d_f += cvec[i] * dx[i];
      for (int r = 0; r < k; ++r) {
        double d = 0.0;
        for (int i = 0; i < n; ++i) d += A[i * k + r] * dx[i];
        const double br = bvec[r];
        d_eq += (br > 0 ? 1.0 : (br < 0 ? -1.0 : 0.0)) * d;
      }
      if (k && null_space) {  // no multipliers: inequality (18.36), nonlinear.cc:492-500
        double l1 = 0.0, quad = 0.0;
        for (int r = 0; r < k; ++r) l1 += fabs(bvec[r]);
        l1 = fmax(l1, 2.220446049250313e-16);
        for (int i = 0; i < n; ++i) {
          double acc = 0.0;
          for (int j = 0; j < n; ++j) acc += (i >= j ? G[j * n + i] : G[i * n + j]) * dx[j];
          quad += dx[i] * acc;
        }
        const double np = (d_f + 0.5 * fmax(0.0, quad)) / ((1 - prm.equality_penalty_rho) * l1);
        if (np > penalty) penalty = np * prm.equality_penalty_scale_factor;
      } else if (k) {
        const double* y = W_.vars + n + m;
        double linf = fabs(y[0]);
        for (int r = 1; r < k; ++r) linf = fmax(linf, fabs(y[r]));
        if (linf > penalty) penalty = linf * prm.equality_penalty_scale_factor;
      }
      // SelectStepSize (:351-416)
      const double dd_total = d_f + penalty * d_eq, dd_linf = fmax(fabs(d_f), fabs(d_eq));
      const double phi0 = pre_f + penalty * pre_eq;
      int step_result = 1;  // 0 SUCCESS 1 MAX_ITERATIONS 2 FIRST_ORDER_SATISFIED 3 POSITIVE_DERIVATIVE 4 NON_FINITE 5 INVALID_ALPHA
      double alpha = 1.0, last_f = 0.0, last_eq = 0.0;
      double a_prev = 0.0, phi_prev = 0.0, a_prev2 = 0.0, phi_prev2 = 0.0;  // last two logged steps
      for (int it = 0; it <= prm.max_line_search_iterations && it < kMaxSteps; ++it) {
        if (prm.line_search_strategy == MO_LS_POLYNOMIAL_APPROXIMATION) {
          if (it > 0) {
            double na = 0.0;
            bool has;
            if (it == 1) {  // QuadraticApproxMinimum, :531-538
              const double numerator = phi_prev - dd_total * a_prev - phi0;
              has = !(dd_total > 0 || numerator <= 0);
              if (has) na = -dd_total * a_prev * a_prev / (2.0 * numerator);
            } else {        // CubicApproxCoeffs + CubicApproxMinimum, :563-605
              const double a0 = a_prev2, a1 = a_prev;
              const double ma = a0 * a0 * a0, mb = a0 * a0, mc = a1 * a1 * a1, md = a1 * a1;
              const double rhs0 = phi_prev2 - phi0 - dd_total * a0, rhs1 = phi_prev - phi0 - dd_total * a1;
              const double invdet = 1.0 / (ma * md - mb * mc);
              const double ca = (md * invdet) * rhs0 + (-mb * invdet) * rhs1;
              const double cb2 = (-mc * invdet) * rhs0 + (ma * invdet) * rhs1;
              const double arg = cb2 * cb2 - 3 * ca * dd_total;
              has = !(ca == 0.0 || arg < -1.0e-12);
              if (has) na = (-cb2 + sqrt(fmax(arg, 0.))) / (3 * ca);
            }
            if (!has || !isfinite(na) || na <= 0.0 || na >= alpha) { step_result = 5; break; }
            alpha = na;
          }
        } else if (it > 0) {
          alpha = alpha * prm.armijo_search_tau;
        }
        for (int i = lane; i < n; i += 32) {  // RetractCandidateVars, :160-168
          const double v = x[i] + dx[i] * alpha;
          cand[i] = prm.retraction == MO_RETRACT_MOD_PI ? mod_pi(v) : v;
        }
        __syncwarp();
        family_errors<F>(C, a, p, cand, rc, last_f, last_eq);
        a_prev2 = a_prev; phi_prev2 = phi_prev;
        a_prev = alpha; phi_prev = last_f + penalty * last_eq;
        ++ls_total;
        if (!isfinite(last_f) || !isfinite(last_eq)) { step_result = 4; break; }
        if (dd_linf < prm.absolute_first_derivative_tol) { step_result = 2; break; }
        else if (dd_total > 0) { step_result = 3; break; }
        if (phi_prev <= phi0 + dd_total * alpha * 1.0e-4) { step_result = 0; break; }
      }
      // UpdateLambdaAndCheckExitConditions (:296-343)
      int exit_state = -1;
      if (step_result == 0) {
        __syncwarp();
        for (int i = lane; i < n; i += 32) x[i] = cand[i];
        __syncwarp();
        lambda = fmax(lambda * (state == 1 ? prm.lambda_decrease_on_restore : prm.lambda_decrease_on_success), prm.min_lambda);
        state = 0;
        if (fmax(last_f, last_eq) < prm.absolute_exit_tol) exit_state = MO_NLS_SATISFIED_ABSOLUTE_TOL;
        else if (phi_prev > phi0 * (1 - prm.relative_exit_tol)) exit_state = MO_NLS_SATISFIED_RELATIVE_TOL;
      } else if (step_result == 2) {
        exit_state = MO_NLS_SATISFIED_FIRST_ORDER_TOL;
      } else if (step_result == 1 || step_result == 3) {
        if (state == 0) { lambda = fmax(prm.lambda_failure_init, lambda * 10.0); state = 1; }
        else lambda *= 10.0;
        if (lambda > prm.max_lambda) exit_state = MO_NLS_MAX_LAMBDA;
      }
      ++iters;
      ++iter;
      if (exit_state >= 0) { term = exit_state; ended = true; }
      else if (iter >= prm.max_iterations) ended = true;  // term stays MAX_ITERATIONS
    }
    IK_T(5);  // phase C (with the wait after B when that barrier is on)
    if (busy && ended) finish();
  }
#ifdef MO_IK_TRACE
  if (blockIdx.x == 0 && threadIdx.x == 0)
    printf("ik trace W=%d: fetch %lld | wait(open) %lld | A linearize %lld | wait(A) %lld | B qp %lld | C line search %lld | finish %lld\n", W,
           s_ik_tr[0], s_ik_tr[1], s_ik_tr[2], s_ik_tr[3], s_ik_tr[4], s_ik_tr[5], s_ik_tr[6]);
#endif
}

template <int W, int MINB>
constexpr auto nls_chain_ik_kernel = nls_solve_kernel<ChainFamily, W, MINB>;

}  // namespace mo
