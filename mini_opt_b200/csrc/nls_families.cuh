// Functor-defined residual families for the device SQP + LM kernel (nls_solve_kernel, nls_chain.cuh): the device-side
// counterpart of the reference's MakeResidual<R, P>(index, functor) (include/mini_opt/residual.hpp:121-143). A problem
// class is ONE struct with compile-time sizes and two device functions; FunctorFamily<Fn> adapts it to the family
// interface and nls_families.cu instantiates the kernel for it:
//
//   struct MyResidual {
//     static constexpr int N = ..., R_COST = ..., COST_BLOCK = ..., R_EQ = ...;   // COST_BLOCK: rows per cost Residual
//     // r[R_COST] and, unless J == nullptr, the column-major R_COST x N Jacobian at x; `data` = this problem's constants
//     static __device__ void cost(const double* x, const double* data, double* r, double* J);
//     static __device__ void equality(const double* x, const double* data, double* r, double* J);  // R_EQ rows
//   };
//
// The three families below are the reference's own nonlinear test problems (test/nonlinear_test.cc:375-387 Rosenbrock,
// :580-594 Himmelblau as two one-row costs, :721-760 sphere with two product equality constraints); they are pinned
// against the unmodified reference start point by start point (tests/golden/nls_host_reference.txt).
// Adding a family: write the struct, add it to MO_NLS_FUNCTOR_FAMILIES in nls_families.cu, rebuild.
#pragma once
#include "nls_chain.cuh"

namespace mo {

struct FunctorArgs {
  int n, n_cost, cost_block, n_eq, m;
  long long batch;
  const int* ineq_var;      // [m] shared by the batch
  const double* ineq_a;     // [m]
  const double* ineq_b;     // [m]
  const double* x0;         // [batch][n]
  const double* data;       // optional [batch][data_stride]
  int data_stride;
};

template <class Fn>
struct FunctorFamily {
  using Args = FunctorArgs;
  struct Work {};
  static_assert(Fn::R_COST <= kMaxChainRows && Fn::R_EQ <= kMaxChainRows, "rows per problem");
  __host__ __device__ static size_t smem_doubles(const Args&) { return 0; }
  __device__ static double* carve(Work&, double* q, const Args&) { return q; }
  __device__ static void setup(Work&, const Args&) {}
  __device__ static void begin(Work&, const Args&, long long) {}
  // The functor is scalar code: lane 0 evaluates it (a residual of a handful of rows has nothing to spread over 32 lanes;
  // the linearization, the QP and the line search logic around it stay warp-parallel).
  __device__ static void cost_rows(Work&, const Args& a, long long p, const double* x, double* r, double* J) {
    if ((threadIdx.x & 31) == 0) Fn::cost(x, a.data ? a.data + p * a.data_stride : nullptr, r, J);
    __syncwarp();
  }
  __device__ static void eq_rows(Work&, const Args& a, long long p, const double* x, double* r, double* J) {
    if ((threadIdx.x & 31) == 0) Fn::equality(x, a.data ? a.data + p * a.data_stride : nullptr, r, J);
    __syncwarp();
  }
};

// h(x, y) = [a - x, sqrt(b) (y - x^2)], a = 1, b = 100 unless data = {a, b}  (nonlinear_test.cc:375-387)
struct RosenbrockResidual {
  static constexpr int N = 2, R_COST = 2, COST_BLOCK = 2, R_EQ = 0;
  static __device__ void cost(const double* x, const double* data, double* r, double* J) {
    const double a = data ? data[0] : 1.0, sb = sqrt(data ? data[1] : 100.0);
    r[0] = a - x[0];
    r[1] = sb * (x[1] - x[0] * x[0]);
    if (J) { J[0] = -1.0; J[1] = -2 * x[0] * sb; J[2] = 0.0; J[3] = sb; }
  }
  static __device__ void equality(const double*, const double*, double*, double*) {}
};

// two one-row costs x^2 + y - 11 and x + y^2 - 7 (nonlinear_test.cc:580-594)
struct HimmelblauResidual {
  static constexpr int N = 2, R_COST = 2, COST_BLOCK = 1, R_EQ = 0;
  static __device__ void cost(const double* x, const double*, double* r, double* J) {
    r[0] = x[0] * x[0] + x[1] - 11;
    r[1] = x[0] + x[1] * x[1] - 7;
    if (J) { J[0] = 2 * x[0]; J[1] = 1.0; J[2] = 1.0; J[3] = 2 * x[1]; }
  }
  static __device__ void equality(const double*, const double*, double*, double*) {}
};

// cost x (identity Jacobian, 6 variables) subject to x0 x1 = 4 and x2 x3 = 9 unless data = {v0, v1}
// (nonlinear_test.cc:721-760); equality-only: the kernel takes the null-space solver like the reference
struct SphereProductResidual {
  static constexpr int N = 6, R_COST = 6, COST_BLOCK = 6, R_EQ = 2;
  static __device__ void cost(const double* x, const double*, double* r, double* J) {
    for (int i = 0; i < N; ++i) r[i] = x[i];
    if (J)
      for (int e = 0; e < N * N; ++e) J[e] = (e % (N + 1)) == 0 ? 1.0 : 0.0;
  }
  static __device__ void equality(const double* x, const double* data, double* r, double* J) {
    r[0] = x[0] * x[1] - (data ? data[0] : 4.0);
    r[1] = x[2] * x[3] - (data ? data[1] : 9.0);
    if (J) {  // 2 x 6 column-major
      for (int e = 0; e < 2 * N; ++e) J[e] = 0.0;
      J[0 * 2 + 0] = x[1]; J[1 * 2 + 0] = x[0];
      J[2 * 2 + 1] = x[3]; J[3 * 2 + 1] = x[2];
    }
  }
};

}  // namespace mo
