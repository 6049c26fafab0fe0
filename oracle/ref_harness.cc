// TEST INFRASTRUCTURE — not part of the product. Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load what this file builds.
//
// Thin extern "C" harness around the UNMODIFIED reference (gareth-cross/mini_opt) so that the same
// byte buffers that go to the CUDA path can be solved by the reference's own
// QPInteriorPointSolver (source/qp.cc) on the host. It is compiled together with the reference's
// sources where they lie under /root/reference by oracle/Makefile into oracle/_ref/libminiopt_ref.so.
// Nothing from the reference is copied into this repository.
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include <Eigen/Dense>

#include "mini_opt/qp.hpp"

#include "ref_friend.h"

#include "mini_opt_b200.h"


using namespace mini_opt;

namespace {

// Unpack problem `p` of the flat batch into a mini_opt::QP.
void UnpackQP(const mo_qp_batch& b, int64_t p, QP* qp) {
  const int n = b.n, k = b.k, m = b.m;
  qp->G = Eigen::Map<const Eigen::MatrixXd>(b.G + p * n * n, n, n);
  qp->c = Eigen::Map<const Eigen::VectorXd>(b.c + p * n, n);
  if (k > 0) {
    qp->A_eq = Eigen::Map<const Eigen::MatrixXd>(b.A_eq + p * k * n, k, n);
    qp->b_eq = Eigen::Map<const Eigen::VectorXd>(b.b_eq + p * k, k);
  } else {
    qp->A_eq.resize(0, 0);
    qp->b_eq.resize(0);
  }
  qp->constraints.clear();
  const int count = b.ineq_count ? b.ineq_count[p] : m;
  for (int i = 0; i < count; ++i) {
    if (b.flags & MO_QP_FLAG_BOX_PAIRS) {
      qp->constraints.emplace_back(i / 2, (i % 2 == 0) ? 1.0 : -1.0, b.ineq_b[p * m + i]);
    } else {
      qp->constraints.emplace_back(b.ineq_var[p * m + i], b.ineq_a[p * m + i], b.ineq_b[p * m + i]);
    }
  }
}

QPInteriorPointSolver::Params ToParams(const mo_qp_params& q) {
  QPInteriorPointSolver::Params p{};
  p.initial_mu = q.initial_mu;
  p.sigma = q.sigma;
  p.termination_kkt_tol = q.termination_kkt_tol;
  p.termination_complementarity_tol = q.termination_complementarity_tol;
  p.max_iterations = q.max_iterations;
  p.barrier_strategy = static_cast<BarrierStrategy>(q.barrier_strategy);
  p.decrease_mu_only_on_small_error = q.decrease_mu_only_on_small_error != 0;
  p.initial_guess_method = static_cast<InitialGuessMethod>(q.initial_guess_method);
  p.initialize_mu_with_complementarity = q.initialize_mu_with_complementarity != 0;
  return p;
}

void WriteKKT(const KKTError& e, double* out) {
  out[0] = e.r_dual;
  out[1] = e.r_comp;
  out[2] = e.r_primal_eq;
  out[3] = e.r_primal_ineq;
}

// Scatter the solver's [x|s|y|z] (sized for `count` constraints) into the fixed-stride slot sized for m.
void StoreVars(const mo_qp_batch& b, int count, const Eigen::VectorXd& v, double* dst) {
  const int n = b.n, k = b.k, m = b.m;
  const double nan = std::numeric_limits<double>::quiet_NaN();
  for (int i = 0; i < n; ++i) dst[i] = v[i];
  for (int i = 0; i < m; ++i) dst[n + i] = i < count ? v[n + i] : nan;
  for (int i = 0; i < k; ++i) dst[n + m + i] = v[n + count + i];
  for (int i = 0; i < m; ++i) dst[n + m + k + i] = i < count ? v[n + count + k + i] : nan;
}

void LoadVars(const mo_qp_batch& b, int count, const double* src, Eigen::VectorXd* v) {
  const int n = b.n, k = b.k, m = b.m;
  v->resize(n + 2 * count + k);
  for (int i = 0; i < n; ++i) (*v)[i] = src[i];
  for (int i = 0; i < count; ++i) (*v)[n + i] = src[n + i];
  for (int i = 0; i < k; ++i) (*v)[n + count + i] = src[n + m + i];
  for (int i = 0; i < count; ++i) (*v)[n + count + k + i] = src[n + m + k + i];
}

// `prepared`: problems [0, batch) already unpacked (the timed entry point keeps UnpackQP outside its clock,
// BASELINE.md section 4); nullptr: unpack on the fly.
void SolveRange(const mo_qp_batch& b, const mo_qp_params& params, const mo_qp_out& out, int64_t begin,
                int64_t end, std::atomic<int64_t>* total_iters, const std::vector<QP>* prepared = nullptr) {
  const auto ref_params = ToParams(params);
  const int64_t stride = b.n + 2 * b.m + b.k;
  QP local;
  QPInteriorPointSolver solver;  // re-used across problems like test/qp_test.cc:531-539
  int64_t iters_sum = 0;
  for (int64_t p = begin; p < end; ++p) {
    if (!prepared) UnpackQP(b, p, &local);
    const QP& qp = prepared ? (*prepared)[p] : local;
    const int count = static_cast<int>(qp.constraints.size());
    int status = -1;
    std::size_t iters = 0;
    KKTError last{};
    const double nan = std::numeric_limits<double>::quiet_NaN();
    double lag[2] = {nan, nan};
    try {
      solver.Setup(&qp);
      if (ref_params.initial_guess_method == InitialGuessMethod::USER_PROVIDED) {
        Eigen::VectorXd v;
        LoadVars(b, count, out.vars + p * stride, &v);
        solver.SetVariables(v);
      }
      const QPInteriorPointSolverOutputs o = solver.Solve(ref_params);
      status = static_cast<int>(o.termination_state);
      iters = o.iterations.size();
      if (!o.iterations.empty()) last = o.iterations.back().kkt_final;
      if (o.lagrange_multipliers) {
        lag[0] = o.lagrange_multipliers->min;
        lag[1] = o.lagrange_multipliers->l_infinity;
      }
      if (out.iter_log) {
        double* log = out.iter_log + p * params.max_iterations * MO_QP_ITER_LOG_STRIDE;
        for (std::size_t it = 0; it < iters; ++it) {
          const auto& I = o.iterations[it];
          double* row = log + it * MO_QP_ITER_LOG_STRIDE;
          WriteKKT(I.kkt_initial, row);
          WriteKKT(I.kkt_final, row + 4);
          row[8] = I.ip_outputs.mu;
          row[9] = I.ip_outputs.alpha.primal;
          row[10] = I.ip_outputs.alpha.dual;
          row[11] = I.ip_outputs.alpha_probe.primal;
          row[12] = I.ip_outputs.alpha_probe.dual;
          row[13] = I.ip_outputs.mu_affine;
        }
      }
    } catch (const FailedFactorization&) {
      status = MO_QP_FAILED_FACTORIZATION;
    } catch (const std::exception&) {
      // F_ASSERT(all s > 0), qp.cc:285 (assert::default_error derives from std::exception)
      status = MO_QP_NONPOSITIVE_SLACK;
    }
    StoreVars(b, count, solver.variables(), out.vars + p * stride);
    out.status[p] = status;
    out.iters[p] = static_cast<int32_t>(iters);
    if (out.kkt_final) WriteKKT(last, out.kkt_final + p * 4);
    if (out.lagrange) {
      out.lagrange[p * 2] = lag[0];
      out.lagrange[p * 2 + 1] = lag[1];
    }
    iters_sum += static_cast<int64_t>(iters);
  }
  if (total_iters) total_iters->fetch_add(iters_sum);
}

}  // namespace

extern "C" {

// Reference QPInteriorPointSolver::Solve on every problem of a host-resident batch, one solver
// instance per std::thread over a static contiguous partition. Returns the total
// number of IP iterations (Newton steps), or a negative mo_error.
int64_t ref_qp_solve_batch(const mo_qp_batch* b, const mo_qp_params* params, const mo_qp_out* out,
                           int nthreads) {
  if (!b || !params || !out || !out->vars || !out->status || !out->iters) return MO_ERR_INVALID_ARGUMENT;
  if (nthreads < 1) nthreads = 1;
  std::atomic<int64_t> total{0};
  if (nthreads == 1 || b->batch < 2 * nthreads) {
    SolveRange(*b, *params, *out, 0, b->batch, &total);
    return total.load();
  }
  std::vector<std::thread> pool;
  const int64_t chunk = (b->batch + nthreads - 1) / nthreads;
  for (int t = 0; t < nthreads; ++t) {
    const int64_t lo = t * chunk, hi = std::min<int64_t>(b->batch, lo + chunk);
    if (lo >= hi) break;
    pool.emplace_back(SolveRange, std::cref(*b), std::cref(*params), std::cref(*out), lo, hi, &total, nullptr);
  }
  for (auto& th : pool) th.join();
  return total.load();
}

// Same, for timing: every problem is unpacked into a mini_opt::QP first (outside the clock), then the threaded solve
// loop alone is timed with std::chrono; *seconds_out receives the wall time of that loop.
int64_t ref_qp_solve_batch_timed(const mo_qp_batch* b, const mo_qp_params* params, const mo_qp_out* out,
                                 int nthreads, double* seconds_out) {
  if (!b || !params || !out || !out->vars || !out->status || !out->iters || !seconds_out) return MO_ERR_INVALID_ARGUMENT;
  if (nthreads < 1) nthreads = 1;
  std::vector<QP> prepared(static_cast<std::size_t>(b->batch));
  for (int64_t p = 0; p < b->batch; ++p) UnpackQP(*b, p, &prepared[p]);
  std::atomic<int64_t> total{0};
  std::vector<std::thread> pool;
  const int64_t chunk = (b->batch + nthreads - 1) / nthreads;
  const auto t0 = std::chrono::steady_clock::now();
  for (int t = 0; t < nthreads; ++t) {
    const int64_t lo = t * chunk, hi = std::min<int64_t>(b->batch, lo + chunk);
    if (lo >= hi) break;
    pool.emplace_back(SolveRange, std::cref(*b), std::cref(*params), std::cref(*out), lo, hi, &total, &prepared);
  }
  for (auto& th : pool) th.join();
  *seconds_out = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return total.load();
}

// One reference Iterate(mu, strategy) (source/qp.cc:153-201) per problem on caller state.
int ref_qp_newton_step_batch(const mo_qp_batch* b, double mu, int32_t strategy, double* vars,
                             double* delta_out, double* ip_out, double* kkt_out, int32_t* status) {
  if (!b || !vars || !status) return MO_ERR_INVALID_ARGUMENT;
  const int64_t stride = b->n + 2 * b->m + b->k;
  QP qp;
  for (int64_t p = 0; p < b->batch; ++p) {
    UnpackQP(*b, p, &qp);
    const int count = static_cast<int>(qp.constraints.size());
    QPInteriorPointSolver solver(&qp);
    Eigen::VectorXd v;
    LoadVars(*b, count, vars + p * stride, &v);
    solver.SetVariables(v);
    try {
      const IPIterationOutputs o =
          ConstrainedNLSTest::Iterate(solver, mu, static_cast<BarrierStrategy>(strategy));
      status[p] = 0;
      StoreVars(*b, count, solver.variables(), vars + p * stride);
      if (delta_out) StoreVars(*b, count, ConstrainedNLSTest::Delta(solver), delta_out + p * stride);
      if (ip_out) {
        double* r = ip_out + p * 6;
        r[0] = o.mu;
        r[1] = o.alpha.primal;
        r[2] = o.alpha.dual;
        r[3] = o.alpha_probe.primal;
        r[4] = o.alpha_probe.dual;
        r[5] = o.mu_affine;
      }
      if (kkt_out) {
        ConstrainedNLSTest::EvaluateKKT(solver);
        WriteKKT(ConstrainedNLSTest::Errors(solver, mu), kkt_out + p * 4);
      }
    } catch (const FailedFactorization&) {
      status[p] = MO_QP_FAILED_FACTORIZATION;
    } catch (const std::exception&) {
      status[p] = MO_QP_NONPOSITIVE_SLACK;
    }
  }
  return MO_OK;
}

// Reference QPNullSpaceSolver::Solve (source/qp.cc:679-729) per problem: x [batch][n], status [batch]
// (QPNullSpaceTerminationState: 0 SUCCESS, 1 NOT_POSITIVE_DEFINITE). Inequalities of the batch are ignored.
int ref_qp_nullspace_batch(const mo_qp_batch* b, double* x, int32_t* status) {
  if (!b || !x || !status || b->k <= 0) return MO_ERR_INVALID_ARGUMENT;
  QP qp;
  QPNullSpaceSolver solver;
  for (int64_t p = 0; p < b->batch; ++p) {
    UnpackQP(*b, p, &qp);
    qp.constraints.clear();
    const QPNullSpaceTerminationState term = solver.Solve(qp);
    status[p] = static_cast<int32_t>(term);
    for (int i = 0; i < b->n; ++i) x[p * b->n + i] = term == QPNullSpaceTerminationState::SUCCESS ? solver.variables()[i] : 0.0;
  }
  return MO_OK;
}

int ref_hardware_concurrency() { return static_cast<int>(std::thread::hardware_concurrency()); }

}  // extern "C"
