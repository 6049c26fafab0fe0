// TEST INFRASTRUCTURE — not part of the product (see ref_harness.cc).
//
// extern "C" harness around the UNMODIFIED reference's ConstrainedNonlinearLeastSquares
// (source/nonlinear.cc), Residual (include/mini_opt/residual.hpp) and ActuatorChain
// (test/transform_chains.cc), consuming the same flat buffers as the CUDA entry points
// mo_nls_linearize_batch / mo_chain_fk_batch / mo_nls_solve_chain_ik_batch. Host pointers everywhere.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <limits>
#include <memory>
#include <numeric>
#include <thread>
#include <vector>

#include <Eigen/Dense>

#include "geometry_utils/angle_utils.hpp"
#include "mini_opt/nonlinear.hpp"
#include "mini_opt/residual.hpp"
#include "transform_chains.hpp"

#include "mini_opt_b200.h"
#include "ref_friend.h"

using namespace mini_opt;
using Eigen::Dynamic;
using Eigen::Matrix;
using Eigen::VectorXd;

namespace {

std::vector<int> Iota(int n) {
  std::vector<int> v(n);
  std::iota(v.begin(), v.end(), 0);
  return v;
}

// A cost / equality Residual of R rows whose value and Jacobian are read from dense buffers
// (`J` column-major rows x n with leading dimension `ld`, rows [row0, row0+R)).
template <int R>
Residual DenseResidual(int n, const double* const* J, const double* const* r, int ld, int row0) {
  auto fn = [=](const VectorXd&, Matrix<double, R, Dynamic>* const J_out) -> Matrix<double, R, 1> {
    Matrix<double, R, 1> out;
    for (int i = 0; i < R; ++i) out[i] = (*r)[row0 + i];
    if (J_out) {
      for (int c = 0; c < n; ++c)
        for (int i = 0; i < R; ++i) (*J_out)(i, c) = (*J)[c * ld + row0 + i];
    }
    return out;
  };
  return MakeResidual<R, Dynamic>(Iota(n), fn);
}

Residual DenseResidualDyn(int R, int n, const double* const* J, const double* const* r, int ld, int row0) {
  switch (R) {
    case 1: return DenseResidual<1>(n, J, r, ld, row0);
    case 2: return DenseResidual<2>(n, J, r, ld, row0);
    case 3: return DenseResidual<3>(n, J, r, ld, row0);
    case 4: return DenseResidual<4>(n, J, r, ld, row0);
    case 6: return DenseResidual<6>(n, J, r, ld, row0);
    case 8: return DenseResidual<8>(n, J, r, ld, row0);
    default: throw std::runtime_error("unsupported residual block size");
  }
}

// Residual rows w_i . root_t_effector - target_i over one ActuatorChain (the y_res / x_eq pattern of
// test/nonlinear_test.cc:840-859).
template <int R>
Residual ChainResidual(int n, ActuatorChain* chain, const double* w, const double* const* target, int row0) {
  auto fn = [=](const VectorXd& params, Matrix<double, R, Dynamic>* const J_out) -> Matrix<double, R, 1> {
    chain->Update(params);
    const Eigen::Vector3d t = chain->translation();
    Matrix<double, R, 1> out;
    for (int i = 0; i < R; ++i) {
      const Eigen::Map<const Eigen::Vector3d> wi(w + 3 * (row0 + i));
      out[i] = wi.dot(t) - (*target)[row0 + i];
    }
    if (J_out) {
      J_out->resize(R, n);
      for (int i = 0; i < R; ++i) {
        const Eigen::Map<const Eigen::Vector3d> wi(w + 3 * (row0 + i));
        J_out->row(i) = wi.transpose() * chain->translation_D_params();
      }
    }
    return out;
  };
  return MakeResidual<R, Dynamic>(Iota(n), fn);
}

Residual ChainResidualDyn(int R, int n, ActuatorChain* chain, const double* w, const double* const* target,
                          int row0) {
  switch (R) {
    case 1: return ChainResidual<1>(n, chain, w, target, row0);
    case 2: return ChainResidual<2>(n, chain, w, target, row0);
    case 3: return ChainResidual<3>(n, chain, w, target, row0);
    default: throw std::runtime_error("unsupported residual block size");
  }
}

void BuildChain(const mo_chain_ik_batch& b, ActuatorChain* chain) {
  chain->links.clear();
  for (int i = 0; i < b.links; ++i) {
    const double* p = b.link_pose + 7 * i;
    std::array<uint8_t, 6> mask{};
    for (int j = 0; j < 6; ++j) mask[j] = b.link_mask[6 * i + j];
    chain->links.emplace_back(Pose(Eigen::Quaterniond(p[0], p[1], p[2], p[3]), Eigen::Vector3d(p[4], p[5], p[6])),
                              mask);
  }
}

ConstrainedNonlinearLeastSquares::Params ToParams(const mo_nls_params& q) {
  ConstrainedNonlinearLeastSquares::Params p{};
  p.max_iterations = q.max_iterations;
  p.max_qp_iterations = q.max_qp_iterations;
  p.termination_kkt_tolerance = q.termination_kkt_tolerance;
  p.absolute_exit_tol = q.absolute_exit_tol;
  p.relative_exit_tol = q.relative_exit_tol;
  p.absolute_first_derivative_tol = q.absolute_first_derivative_tol;
  p.max_line_search_iterations = q.max_line_search_iterations;
  p.line_search_strategy = static_cast<LineSearchStrategy>(q.line_search_strategy);
  p.armijo_search_tau = q.armijo_search_tau;
  p.equality_penalty_initial = q.equality_penalty_initial;
  p.equality_penalty_scale_factor = q.equality_penalty_scale_factor;
  p.equality_penalty_rho = q.equality_penalty_rho;
  p.lambda_initial = q.lambda_initial;
  p.lambda_failure_init = q.lambda_failure_init;
  p.lambda_decrease_on_success = q.lambda_decrease_on_success;
  p.lambda_decrease_on_restore = q.lambda_decrease_on_restore;
  p.max_lambda = q.max_lambda;
  p.min_lambda = q.min_lambda;
  return p;
}

void SolveChainRange(const mo_chain_ik_batch& b, const mo_nls_params& prm, const mo_nls_out& out, int64_t begin,
                     int64_t end, std::atomic<int64_t>* total_qp) {
  const int n = b.n;
  auto chain = std::make_unique<ActuatorChain>();  // one chain per thread: Update() caches (transform_chains.cc:248-258)
  BuildChain(b, chain.get());
  const double* cost_target = nullptr;
  const double* eq_target = nullptr;
  Problem problem{};
  problem.dimension = n;
  const int cb = b.cost_block > 0 ? b.cost_block : b.n_cost;
  for (int row = 0; row < b.n_cost; row += cb)
    problem.costs.push_back(ChainResidualDyn(cb, n, chain.get(), b.cost_w, &cost_target, row));
  for (int row = 0; row < b.n_eq; ++row)
    problem.equality_constraints.push_back(ChainResidualDyn(1, n, chain.get(), b.eq_w, &eq_target, row));
  for (int i = 0; i < b.m; ++i)
    problem.inequality_constraints.emplace_back(b.ineq_var[i], b.ineq_a[i], b.ineq_b[i]);
  ConstrainedNonlinearLeastSquares::Retraction retraction = nullptr;
  if (prm.retraction == MO_RETRACT_MOD_PI) {  // test/nonlinear_test.cc:864-870
    retraction = [](Eigen::VectorXd& x, const ConstVectorBlock& dx, const double alpha) {
      for (int i = 0; i < x.rows(); ++i) x[i] = math::ModPi(x[i] + dx[i] * alpha);
    };
  }
  ConstrainedNonlinearLeastSquares nls(&problem, retraction);
  const auto params = ToParams(prm);
  int64_t qp_sum = 0;
  for (int64_t p = begin; p < end; ++p) {
    cost_target = b.cost_target + p * b.n_cost;
    eq_target = b.n_eq ? b.eq_target + p * b.n_eq : nullptr;
    const VectorXd x0 = Eigen::Map<const VectorXd>(b.x0 + p * n, n);
    int term = MO_NLS_QP_FAILED, iters = 0, qp_iters = 0, ls = 0;
    try {
      const NLSSolverOutputs o = nls.Solve(params, x0);
      term = static_cast<int>(o.termination_state);
      iters = static_cast<int>(o.iterations.size());
      qp_iters = static_cast<int>(o.NumQPIterations());
      ls = static_cast<int>(o.NumLineSearchSteps());
    } catch (const std::exception&) {
      term = MO_NLS_QP_FAILED;
    }
    const VectorXd& x = nls.variables();
    for (int i = 0; i < n; ++i) out.x[p * n + i] = x[i];
    out.termination[p] = term;
    out.iterations[p] = iters;
    if (out.qp_iterations) out.qp_iterations[p] = qp_iters;
    if (out.ls_steps) out.ls_steps[p] = ls;
    if (out.errors) {
      const Errors e = nls.EvaluateNonlinearErrors(x);
      out.errors[2 * p] = e.f;
      out.errors[2 * p + 1] = e.equality;
    }
    qp_sum += qp_iters;
  }
  if (total_qp) total_qp->fetch_add(qp_sum);
}

}  // namespace

extern "C" {

// LinearizeAndFillQP (source/nonlinear.cc:170-214) on dense per-problem Jacobians, through real Residual objects.
int ref_nls_linearize_batch(const mo_nls_linearize_args* a) {
  if (!a) return MO_ERR_INVALID_ARGUMENT;
  const int n = a->n, k = a->r_eq, m = a->m;
  const int cb = a->cost_block > 0 ? a->cost_block : a->r_cost;
  const double *Jc = nullptr, *rc = nullptr, *Je = nullptr, *re = nullptr;
  try {
    Problem problem{};
    problem.dimension = n;
    for (int row = 0; row < a->r_cost; row += cb)
      problem.costs.push_back(DenseResidualDyn(cb, n, &Jc, &rc, a->r_cost, row));
    for (int row = 0; row < k; ++row)
      problem.equality_constraints.push_back(DenseResidualDyn(1, n, &Je, &re, k, row));
    QP qp;
    qp.G.resize(n, n);
    qp.c.resize(n);
    qp.A_eq.resize(k, n);
    qp.b_eq.resize(k);
    for (int64_t p = 0; p < a->batch; ++p) {
      Jc = a->J_cost + p * a->r_cost * n;
      rc = a->r_cost_v + p * a->r_cost;
      Je = k ? a->J_eq + p * k * n : nullptr;
      re = k ? a->r_eq_v + p * k : nullptr;
      problem.inequality_constraints.clear();
      for (int i = 0; i < m; ++i)
        problem.inequality_constraints.emplace_back(a->ineq_var[p * m + i], a->ineq_a[p * m + i], a->ineq_b[p * m + i]);
      const VectorXd x = Eigen::Map<const VectorXd>(a->x + p * n, n);
      const Errors e = ConstrainedNLSTest::Linearize(x, a->lambda[p], problem, &qp);
      std::copy_n(qp.G.data(), n * n, a->G + p * n * n);
      std::copy_n(qp.c.data(), n, a->c + p * n);
      if (k) {
        std::copy_n(qp.A_eq.data(), k * n, a->A_eq + p * k * n);
        std::copy_n(qp.b_eq.data(), k, a->b_eq + p * k);
      }
      for (int i = 0; i < m; ++i) a->ineq_b_shifted[p * m + i] = qp.constraints[i].b;
      a->errors[2 * p] = e.f;
      a->errors[2 * p + 1] = e.equality;
    }
  } catch (const std::exception&) {
    return MO_ERR_INVALID_ARGUMENT;
  }
  return MO_OK;
}

// ActuatorChain::Update (test/transform_chains.cc:165-244) per parameter vector; a fresh chain per problem so that
// the parameter cache never short-circuits.
int ref_chain_fk_batch(const mo_chain_ik_batch* b, const double* params, double* translation, double* rotation_q,
                       double* t_D_params, double* r_D_params) {
  if (!b || !params) return MO_ERR_INVALID_ARGUMENT;
  const int n = b->n;
  try {
    for (int64_t p = 0; p < b->batch; ++p) {
      ActuatorChain chain;
      BuildChain(*b, &chain);
      chain.Update(Eigen::Map<const VectorXd>(params + p * n, n));
      const Eigen::Vector3d t = chain.translation();
      const Eigen::Quaterniond& q = chain.rotation();
      if (translation) std::copy_n(t.data(), 3, translation + 3 * p);
      if (rotation_q) {
        rotation_q[4 * p] = q.w(); rotation_q[4 * p + 1] = q.x(); rotation_q[4 * p + 2] = q.y(); rotation_q[4 * p + 3] = q.z();
      }
      if (t_D_params) std::copy_n(chain.translation_D_params().data(), 3 * n, t_D_params + 3 * n * p);
      if (r_D_params) std::copy_n(chain.rotation_D_params().data(), 3 * n, r_D_params + 3 * n * p);
    }
  } catch (const std::exception&) {
    return MO_ERR_INVALID_ARGUMENT;
  }
  return MO_OK;
}

// ConstrainedNonlinearLeastSquares::Solve per problem, one solver + chain per std::thread. Returns the total
// number of interior point iterations, or a negative mo_error.
int64_t ref_nls_solve_chain_ik_batch(const mo_chain_ik_batch* b, const mo_nls_params* prm, const mo_nls_out* out,
                                     int nthreads) {
  if (!b || !prm || !out || !out->x || !out->termination || !out->iterations) return MO_ERR_INVALID_ARGUMENT;
  if (nthreads < 1) nthreads = 1;
  std::atomic<int64_t> total{0};
  try {
    if (nthreads == 1 || b->batch < 2 * nthreads) {
      SolveChainRange(*b, *prm, *out, 0, b->batch, &total);
      return total.load();
    }
    std::vector<std::thread> pool;
    const int64_t chunk = (b->batch + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
      const int64_t lo = t * chunk, hi = std::min<int64_t>(b->batch, lo + chunk);
      if (lo >= hi) break;
      pool.emplace_back(SolveChainRange, std::cref(*b), std::cref(*prm), std::cref(*out), lo, hi, &total);
    }
    for (auto& th : pool) th.join();
  } catch (const std::exception&) {
    return MO_ERR_INVALID_ARGUMENT;
  }
  return total.load();
}

}  // extern "C"
