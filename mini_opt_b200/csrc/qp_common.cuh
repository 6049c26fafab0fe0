// Shared definitions of the batched interior point QP kernels (sm_100a).
//
// Reference path: QPInteriorPointSolver::Solve / Iterate (gareth-cross/mini_opt, source/qp.cc:100-201).
// Nothing here is a translation of that code: the kernels factor the reduced KKT system once per
// Newton step and solve for the one or two right-hand sides they need (no explicit inverse, no
// Eigen), but the arithmetic they perform is the reference's, cited function by function.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "mini_opt_b200.h"

namespace mo {

// Kernel-side copy of mo_qp_batch / mo_qp_params / mo_qp_out (passed by value as kernel arguments).
struct QPArgs {
  int n, k, m, flags;
  long long batch;
  const double* __restrict__ G;
  const double* __restrict__ c;
  const double* __restrict__ A;
  const double* __restrict__ b;
  const int* __restrict__ ivar;
  const double* __restrict__ ia;
  const double* __restrict__ ib;
  const int* __restrict__ icount;
  // params (QPInteriorPointSolver::Params, qp.hpp:134-164)
  double initial_mu, sigma, kkt_tol, comp_tol;
  int max_iterations, strategy, decrease_mu_only_on_small_error, guess, init_mu_with_comp;
  // outputs
  double* __restrict__ vars;
  int* __restrict__ status;
  int* __restrict__ iters;
  double* __restrict__ kkt_final;
  double* __restrict__ lagrange;
  double* __restrict__ iter_log;
};

// Arguments of the single-Newton-step entry point (Iterate, qp.cc:153-201).
struct StepArgs {
  double mu;
  double* __restrict__ delta_out;
  double* __restrict__ ip_out;
  double* __restrict__ kkt_out;
};

// Reciprocal without the IEEE slow path: MUFU.RCP64H seed (rel. error ~2^-23) + two Newton steps.
// Result is within ~1 ulp for normal, finite, non-zero d; callers screen d == 0 / non-finite.
__device__ __forceinline__ double fast_rcp(double d) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
  double e = fma(-d, x, 1.0);
  x = fma(x, e, x);
  e = fma(-d, x, 1.0);
  x = fma(x, e, x);
  return x;
}

__device__ __forceinline__ double qnan() { return __longlong_as_double(0x7ff8000000000000LL); }

}  // namespace mo
