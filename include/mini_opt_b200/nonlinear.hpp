// C++ host API of mini_opt_b200 for the nonlinear least squares driver: the reference's class surface
// (gareth-cross/mini_opt include/mini_opt/residual.hpp:25-143, include/mini_opt/nonlinear.hpp:24-157,
// include/mini_opt/structs.hpp:150-248) with the Residual callbacks kept on the HOST and the QP solves batched on the
// device. Header only; link with -lminiopt_b200.
//
// `ConstrainedNonlinearLeastSquares::SolveBatch` advances B independent problems of identical shape in lockstep: per
// outer iteration every still-active problem is linearized on the host by its own Residual functors
// (LinearizeAndFillQP, source/nonlinear.cc:170-214), all the QPs go to the device in ONE
// QPInteriorPointSolver::SolveBatch call (ComputeStepDirection, :216-259), and the scalar SQP logic of Solve
// (:75-158: ComputeQPCostDerivative :457-482, SelectPenalty :484-502, SelectStepSize :351-416 with both line search
// strategies :418-605, UpdateLambdaAndCheckExitConditions :296-343) runs per problem on the host, where the user's
// functors live. Problems whose residuals are transform chains can run entirely on the device instead
// (mo_nls_solve_chain_ik_batch). Equality-only problems (the reference switches to QPNullSpaceSolver, :83-89) are
// solved by the interior point kernel with M = 0, which Iterate supports (mu = 0).
#pragma once

#include <algorithm>
#include <array>
#include <cmath>
#include <functional>
#include <limits>
#include <memory>
#include <optional>
#include <utility>
#include <variant>

#include "mini_opt_b200/qp.hpp"

namespace mini_opt_b200 {

enum class LineSearchStrategy { ARMIJO_BACKTRACK = 0, POLYNOMIAL_APPROXIMATION = 1 };  // structs.hpp:147-152
enum class OptimizerState { NOMINAL = 0, ATTEMPTING_RESTORE_LM = 1 };                  // structs.hpp:155-158
enum class StepSizeSelectionResult {                                                   // structs.hpp:208-221
  SUCCESS = 0, MAX_ITERATIONS, FIRST_ORDER_SATISFIED, POSITIVE_DERIVATIVE, FAILURE_NON_FINITE_COST, FAILURE_INVALID_ALPHA
};
enum class NLSTerminationState {                                                       // structs.hpp:225-240
  MAX_ITERATIONS, SATISFIED_ABSOLUTE_TOL, SATISFIED_RELATIVE_TOL, SATISFIED_FIRST_ORDER_TOL, MAX_LAMBDA, QP_INDEFINITE,
  USER_CALLBACK,
  // batched stand-in (== MO_NLS_QP_FAILED): the inner QP threw FailedFactorization / hit F_ASSERT(s > 0) for THIS problem;
  // the batch goes on. The batch-of-one Solve() rethrows like the reference.
  QP_FAILED
};

// names as the reference's ostream operators print them (structs.cc:50-120)
inline const char* ToString(OptimizerState s) { return s == OptimizerState::NOMINAL ? "NOMINAL" : "ATTEMPTING_RESTORE_LM"; }
inline const char* ToString(LineSearchStrategy s) {
  return s == LineSearchStrategy::ARMIJO_BACKTRACK ? "ARMIJO_BACKTRACK" : "POLYNOMIAL_APPROXIMATION";
}
inline const char* ToString(StepSizeSelectionResult s) {
  static const char* const names[] = {"SUCCESS", "MAX_ITERATIONS", "FIRST_ORDER_SATISFIED", "POSITIVE_DERIVATIVE",
                                      "FAILURE_NON_FINITE_COST", "FAILURE_INVALID_ALPHA"};
  return names[static_cast<int>(s)];
}
inline const char* ToString(NLSTerminationState s) {
  static const char* const names[] = {"MAX_ITERATIONS", "SATISFIED_ABSOLUTE_TOL", "SATISFIED_RELATIVE_TOL",
                                      "SATISFIED_FIRST_ORDER_TOL", "MAX_LAMBDA", "QP_INDEFINITE", "USER_CALLBACK", "QP_FAILED"};
  return names[static_cast<int>(s)];
}
inline std::ostream& operator<<(std::ostream& os, OptimizerState s) { return os << ToString(s); }
inline std::ostream& operator<<(std::ostream& os, LineSearchStrategy s) { return os << ToString(s); }
inline std::ostream& operator<<(std::ostream& os, StepSizeSelectionResult s) { return os << ToString(s); }
inline std::ostream& operator<<(std::ostream& os, NLSTerminationState s) { return os << ToString(s); }

struct Errors {  // structs.hpp:169-186
  double f{0}, equality{0};
  double Total(double penalty) const { return f + penalty * equality; }
  double LInfinity() const { return std::max(f, equality); }
  bool ContainsInvalidValues() const { return !std::isfinite(f) || !std::isfinite(equality); }
};
struct DirectionalDerivatives {  // structs.hpp:189-199
  double d_f{0}, d_equality{0};
  double Total(double penalty) const { return d_f + penalty * d_equality; }
  double LInfinity() const { return std::max(std::abs(d_f), std::abs(d_equality)); }
};
struct LineSearchStep { double alpha; Errors errors; };  // structs.hpp:202-205

// One iteration of the outer loop (NLSIteration, structs.hpp:251-290). `qp_outputs` is the reference's variant; its
// `iterations` vector is filled when Params::record_qp_iterations is set. qp_eigenvalues is never computed (JSON: null).
struct NLSIteration {
  int iteration;
  OptimizerState optimizer_state;
  double lambda;
  Errors errors_initial;
  QPInteriorPointTerminationState qp_termination_state;
  int qp_iterations;
  DirectionalDerivatives directional_derivatives;
  double penalty;
  StepSizeSelectionResult step_result;
  std::vector<LineSearchStep> line_search_steps;
  std::variant<QPNullSpaceTerminationState, QPInteriorPointSolverOutputs> qp_outputs{};

  // structs.cc:159-206. ANSI colours as the reference: green 112 / red 160 / orange 202, reset "\x1b[0m".
  std::string ToString(bool use_color = false, bool include_qp = false) const {
    using detail::Fmt;
    using detail::Printf;
    auto color = [use_color](int code) {
      return !use_color ? std::string() : (code < 0 ? std::string("\x1b[0m") : "\x1b[38;5;" + std::to_string(code) + "m");
    };
    std::string r = "Iteration # " + std::to_string(iteration) + ", state = " + mini_opt_b200::ToString(optimizer_state) +
                    ", lambda = " + Fmt(lambda) + "\n";
    r += "  f(0): " + Printf("%.16e", errors_initial.f) + ", c(0): " + Printf("%.16e", errors_initial.equality) +
         ", total: " + Printf("%.16e", errors_initial.Total(penalty)) + "\n";
    r += "  penalty = " + Printf("%.16f", penalty) + "\n";
    const QPInteriorPointSolverOutputs* ip = std::get_if<QPInteriorPointSolverOutputs>(&qp_outputs);
    if (ip) {
      r += std::string("  QP-IP: ") + mini_opt_b200::ToString(ip->termination_state) + ", " +
           std::to_string(ip->iterations.empty() ? static_cast<std::size_t>(ip->num_iterations) : ip->iterations.size()) + "\n";
    } else {
      r += std::string("  QP: ") + mini_opt_b200::ToString(std::get<QPNullSpaceTerminationState>(qp_outputs)) + "\n";
    }
    r += "  df/dalpha = " + Fmt(directional_derivatives.d_f) + ", dc/dalpha = " + Fmt(directional_derivatives.d_equality) + "\n";
    r += "  Search result: " + color(step_result == StepSizeSelectionResult::SUCCESS ? 112 : 160) +
         mini_opt_b200::ToString(step_result) + color(-1) + "\n";
    std::size_t i = 0;
    for (const LineSearchStep& step : line_search_steps) {
      const std::string idx = std::to_string(i++);
      r += "  f(" + idx + "): " + Printf("%.16e", step.errors.f) + ", c(" + idx + "): " + Printf("%.16e", step.errors.equality) +
           ", total: " + Printf("%.16e", step.errors.Total(penalty)) + ", alpha = " + Printf("%.10f", step.alpha) + "\n";
    }
    if (ip && include_qp)
      for (const QPInteriorPointIteration& it : ip->iterations) r += it.ToString();
    return r;
  }
  std::string ToJson() const {  // serialization.cc:112-123, keys in nlohmann's alphabetical order
    using detail::Json;
    auto errors = [](const Errors& e) { return "{\"equality\":" + Json(e.equality) + ",\"f\":" + Json(e.f) + "}"; };
    std::string r = "{\"directional_derivatives\":{\"d_equality\":" + Json(directional_derivatives.d_equality) +
                    ",\"d_f\":" + Json(directional_derivatives.d_f) + "},\"errors_initial\":" + errors(errors_initial) +
                    ",\"iteration\":" + std::to_string(iteration) + ",\"lambda\":" + Json(lambda) + ",\"line_search_steps\":[";
    for (std::size_t i = 0; i < line_search_steps.size(); ++i)
      r += std::string(i ? "," : "") + "{\"alpha\":" + Json(line_search_steps[i].alpha) + ",\"errors\":" +
           errors(line_search_steps[i].errors) + "}";
    r += std::string("],\"optimizer_state\":\"") + mini_opt_b200::ToString(optimizer_state) + "\",\"penalty\":" + Json(penalty) +
         ",\"qp_eigenvalues\":null,\"qp_outputs\":";
    if (const QPInteriorPointSolverOutputs* ip = std::get_if<QPInteriorPointSolverOutputs>(&qp_outputs)) r += ip->ToJson();
    else r += std::string("\"") + mini_opt_b200::ToString(std::get<QPNullSpaceTerminationState>(qp_outputs)) + "\"";
    return r + ",\"step_result\":\"" + mini_opt_b200::ToString(step_result) + "\"}";
  }
};
struct NLSSolverOutputs {  // structs.hpp:293-301
  NLSTerminationState termination_state = NLSTerminationState::MAX_ITERATIONS;
  std::vector<NLSIteration> iterations;
  Vector variables;  // batched API: the final variables travel with the outputs
  int qp_failure_status = 0;  // QP_FAILED: MO_QP_FAILED_FACTORIZATION or MO_QP_NONPOSITIVE_SLACK of the inner solve

  std::size_t NumQPIterations() const {  // structs.cc:208-219: a null-space solve counts as one
    std::size_t total = 0;
    for (const NLSIteration& it : iterations) total += static_cast<std::size_t>(it.qp_iterations);
    return total;
  }
  std::size_t NumLineSearchSteps() const {  // structs.cc:221-226
    std::size_t total = 0;
    for (const NLSIteration& it : iterations) total += it.line_search_steps.size();
    return total;
  }
  std::size_t NumFailedLineSearches() const {  // structs.cc:228-237
    std::size_t total = 0;
    for (const NLSIteration& it : iterations)
      total += it.step_result != StepSizeSelectionResult::SUCCESS && it.step_result != StepSizeSelectionResult::FIRST_ORDER_SATISFIED;
    return total;
  }
  std::string ToString(bool use_color = false, bool include_qp = false) const {  // structs.cc:239-268
    std::string r;
    for (const NLSIteration& it : iterations) r += it.ToString(use_color, include_qp);
    int code = -1;
    switch (termination_state) {
      case NLSTerminationState::SATISFIED_ABSOLUTE_TOL:
      case NLSTerminationState::SATISFIED_RELATIVE_TOL:
      case NLSTerminationState::SATISFIED_FIRST_ORDER_TOL: code = 112; break;
      case NLSTerminationState::MAX_LAMBDA:
      case NLSTerminationState::MAX_ITERATIONS:
      case NLSTerminationState::USER_CALLBACK: code = 202; break;
      case NLSTerminationState::QP_INDEFINITE:
      case NLSTerminationState::QP_FAILED: code = 160; break;
    }
    const std::string on = use_color ? "\x1b[38;5;" + std::to_string(code) + "m" : "", off = use_color ? "\x1b[0m" : "";
    return r + "  termination = " + on + mini_opt_b200::ToString(termination_state) + off + "\n";
  }
  std::string ToJson() const {  // serialization.cc:126-134
    std::string r = "{\"iterations\":[";
    for (std::size_t i = 0; i < iterations.size(); ++i) r += (i ? "," : "") + iterations[i].ToJson();
    return r + "],\"termination_state\":\"" + mini_opt_b200::ToString(termination_state) + "\"}";
  }
};

// Type-erased residual (residual.hpp:25-118). `Model<R, P>` wraps a functor
//   std::array<double, R> f(const std::array<double, P>& params, std::array<double, R * P>* J_or_null)
// with J column-major R x P like Eigen::Matrix<double, R, P>.
class Residual {
 public:
  struct Concept {
    virtual ~Concept() = default;
    virtual int Dimension() const = 0;
    virtual const std::vector<int>& Index() const = 0;
    // r[Dimension()], J (Dimension() x Index().size(), column-major) or nullptr
    virtual void Evaluate(const Vector& params, double* r, double* J) const = 0;
  };
  template <int R, int P, class F>
  struct Model final : Concept {
    Model(const std::array<int, P>& index, F f) : index_(index.begin(), index.end()), func_(std::move(f)) {}
    int Dimension() const override { return R; }
    const std::vector<int>& Index() const override { return index_; }
    void Evaluate(const Vector& params, double* r, double* J) const override {
      std::array<double, P> p;  // GatherParams, residual.hpp:151-169
      for (int i = 0; i < P; ++i) {
        if (index_[i] < 0 || index_[i] >= params.rows()) throw std::out_of_range("Index exceeds the number of params");
        p[i] = params[index_[i]];
      }
      std::array<double, R * P> Jl{};
      const std::array<double, R> rv = func_(p, J ? &Jl : nullptr);
      std::copy(rv.begin(), rv.end(), r);
      if (J) std::copy(Jl.begin(), Jl.end(), J);
    }
    std::vector<int> index_;
    F func_;
  };

  Residual() = default;
  explicit Residual(std::shared_ptr<const Concept> impl) : impl_(std::move(impl)) {}
  int Dimension() const { return impl_->Dimension(); }
  const std::vector<int>& GetIndex() const { return impl_->Index(); }

  // residual.hpp:175-180
  void ErrorVector(const Vector& params, double* b_out) const { impl_->Evaluate(params, b_out, nullptr); }
  double QuadraticError(const Vector& params) const {
    std::vector<double> r(Dimension());
    ErrorVector(params, r.data());
    double s = 0;
    for (double v : r) s += v * v;
    return 0.5 * s;
  }
  // residual.hpp:186-226: H(lower) += J'J, b += J'r over the residual's parameter index; returns 1/2 |r|^2
  double UpdateHessian(const Vector& params, Matrix* H, Vector* b) const {
    const int R = Dimension(), N = static_cast<int>(GetIndex().size());
    std::vector<double> r(R), J(static_cast<std::size_t>(R) * N);
    impl_->Evaluate(params, r.data(), J.data());
    for (int row_local = 0; row_local < N; ++row_local) {
      const int row_global = GetIndex()[row_local];
      for (int col_local = 0; col_local <= row_local; ++col_local) {
        const int col_global = GetIndex()[col_local];
        double JtT = 0;
        for (int e = 0; e < R; ++e) JtT += J[row_local * R + e] * J[col_local * R + e];
        if (col_global <= row_global) (*H)(row_global, col_global) += JtT;
        else (*H)(col_global, row_global) += JtT;
      }
      double Jtr = 0;
      for (int e = 0; e < R; ++e) Jtr += J[row_local * R + e] * r[e];
      (*b)[row_global] += Jtr;
    }
    double s = 0;
    for (double v : r) s += v * v;
    return 0.5 * s;
  }
  // residual.hpp:230-250: rows [row, row + Dimension()) of A <- J scattered by index, of b <- r
  void UpdateJacobian(const Vector& params, Matrix* A, Vector* b, int row) const {
    const int R = Dimension(), N = static_cast<int>(GetIndex().size());
    std::vector<double> r(R), J(static_cast<std::size_t>(R) * N);
    impl_->Evaluate(params, r.data(), J.data());
    for (int e = 0; e < R; ++e) (*b)[row + e] = r[e];
    for (int col_local = 0; col_local < N; ++col_local)
      for (int e = 0; e < R; ++e) (*A)(row + e, GetIndex()[col_local]) = J[col_local * R + e];
  }

 private:
  std::shared_ptr<const Concept> impl_;
};

// residual.hpp:121-143
template <int R, int P, class F>
Residual MakeResidual(const std::array<int, P>& index, F&& f) {
  return Residual(std::make_shared<Residual::Model<R, P, std::decay_t<F>>>(index, std::forward<F>(f)));
}

#ifdef MINI_OPT_B200_WITH_EIGEN
}  // namespace mini_opt_b200
#include <Eigen/Core>
namespace mini_opt_b200 {
// With Eigen on the include path (-DMINI_OPT_B200_WITH_EIGEN): residual functors with the REFERENCE's signature
// (residual.hpp:121-143)   Eigen::Matrix<double, R, 1> f(const Eigen::Matrix<double, P, 1>&, Eigen::Matrix<double, R, P>* J)
// are accepted as they are: a reference user's MakeResidual<R, P>(index, functor) call keeps compiling.
template <int R, int P, class F>
Residual MakeResidualEigen(const std::array<int, P>& index, F f) {
  return MakeResidual<R, P>(index, [f](const std::array<double, P>& params, std::array<double, R * P>* J) {
    Eigen::Matrix<double, P, 1> x;
    for (int i = 0; i < P; ++i) x[i] = params[i];
    Eigen::Matrix<double, R, P> Jm;
    const Eigen::Matrix<double, R, 1> r = f(x, J ? &Jm : nullptr);
    if (J)
      for (int c = 0; c < P; ++c)
        for (int i = 0; i < R; ++i) (*J)[c * R + i] = Jm(i, c);  // column-major, like Eigen's default
    std::array<double, R> out;
    for (int i = 0; i < R; ++i) out[i] = r[i];
    return out;
  });
}
inline Eigen::VectorXd ToEigen(const Vector& v) { return Eigen::Map<const Eigen::VectorXd>(v.data(), static_cast<Eigen::Index>(v.size())); }
inline Vector FromEigen(const Eigen::VectorXd& v) { return Vector(v.data(), v.data() + v.size()); }
#endif

// nonlinear.hpp:24-35: min sum 1/2 |f_i(x)|^2  s.t.  c_j(x) = 0,  a x[v] + b >= 0
struct Problem {
  int dimension{0};
  std::vector<Residual> costs;
  std::vector<LinearInequalityConstraint> inequality_constraints;
  std::vector<Residual> equality_constraints;
  void clear() { costs.clear(); inequality_constraints.clear(); equality_constraints.clear(); }
};

class ConstrainedNonlinearLeastSquares {
 public:
  struct Params {  // nonlinear.hpp:37-124, same defaults
    int max_iterations{10};
    int max_qp_iterations{10};
    double termination_kkt_tolerance{1.0e-6};
    double absolute_exit_tol{1.0e-12};
    double relative_exit_tol{1.0e-5};
    double absolute_first_derivative_tol{1.0e-6};
    int max_line_search_iterations{2};
    LineSearchStrategy line_search_strategy{LineSearchStrategy::POLYNOMIAL_APPROXIMATION};
    double armijo_search_tau{0.8};
    double equality_penalty_initial{1.0};
    double equality_penalty_scale_factor{1.01};
    double equality_penalty_rho{0.1};
    double lambda_initial{0.0};
    double lambda_failure_init{1.0e-2};
    double lambda_decrease_on_success{0.1};
    double lambda_decrease_on_restore{0.8};
    double max_lambda{1.};
    double min_lambda{0.};
    // batched API only: keep the per-step log of every inner QP solve in NLSIteration::qp_outputs.iterations (the
    // reference always does; here it costs 112 bytes per Newton step of device-to-host traffic)
    bool record_qp_iterations{false};
  };
  // nonlinear.hpp:127: x <- x (+) alpha * dx
  using Retraction = std::function<void(Vector& x, const Vector& dx, double alpha)>;
  // nonlinear.hpp:150-153 / nonlinear.cc:145-151: called after every outer iteration; returning false ends the problem with
  // USER_CALLBACK unless that iteration terminated it anyway. Batched form: (problem index in the batch, the iteration).
  using ExitCallback = std::function<bool(const NLSIteration&)>;
  using BatchExitCallback = std::function<bool(std::size_t, const NLSIteration&)>;

  explicit ConstrainedNonlinearLeastSquares(const Problem* problem, Retraction retraction = nullptr)
      : p_(problem), custom_retraction_(std::move(retraction)) {
    if (!p_) throw std::invalid_argument("Must have a valid problem");
  }

  // One problem = a batch of one.
  NLSSolverOutputs Solve(const Params& params, const Vector& variables, int device = 0) {
    BatchExitCallback cb;
    if (user_exit_callback_) cb = [this](std::size_t, const NLSIteration& it) { return user_exit_callback_(it); };
    std::vector<NLSSolverOutputs> out = SolveBatch({p_}, params, {variables}, custom_retraction_, device, cb);
    variables_ = out[0].variables;
    if (out[0].termination_state == NLSTerminationState::QP_FAILED) {  // one problem: throw like the reference
      if (out[0].qp_failure_status == MO_QP_NONPOSITIVE_SLACK) throw std::runtime_error("Some slack variables s <= 0");
      throw FailedFactorization("Failed to solve self-adjoint system");
    }
    return std::move(out[0]);
  }
  const Vector& variables() const { return variables_; }
  template <typename T>
  void SetUserExitCallback(T&& cb) { user_exit_callback_ = std::forward<T>(cb); }

  // nonlinear.cc:279-293
  static Errors EvaluateNonlinearErrors(const Problem& problem, const Vector& vars) {
    Errors e{};
    std::vector<double> buf;
    for (const Residual& cost : problem.costs) {
      buf.assign(cost.Dimension(), 0.0);
      cost.ErrorVector(vars, buf.data());
      double s = 0;
      for (double v : buf) s += v * v;
      e.f += 0.5 * s;
    }
    for (const Residual& eq : problem.equality_constraints) {
      buf.assign(eq.Dimension(), 0.0);
      eq.ErrorVector(vars, buf.data());
      for (double v : buf) e.equality += std::abs(v);
    }
    return e;
  }
  Errors EvaluateNonlinearErrors(const Vector& vars) const { return EvaluateNonlinearErrors(*p_, vars); }

  // nonlinear.cc:170-214
  static Errors LinearizeAndFillQP(const Vector& variables, double lambda, const Problem& problem, QP* qp) {
    const int n = problem.dimension;
    int rows = 0;
    for (const Residual& eq : problem.equality_constraints) rows += eq.Dimension();
    qp->G.resize(n, n);
    qp->c.assign(static_cast<std::size_t>(n), 0.0);
    qp->A_eq.resize(rows, n);
    qp->b_eq.assign(static_cast<std::size_t>(rows), 0.0);
    Errors e{};
    for (const Residual& cost : problem.costs) e.f += cost.UpdateHessian(variables, &qp->G, &qp->c);
    if (lambda > 0)
      for (int i = 0; i < n; ++i) qp->G(i, i) += lambda;
    int row = 0;
    for (const Residual& eq : problem.equality_constraints) {
      eq.UpdateJacobian(variables, &qp->A_eq, &qp->b_eq, row);
      for (int i = 0; i < eq.Dimension(); ++i) e.equality += std::abs(qp->b_eq[row + i]);
      row += eq.Dimension();
    }
    qp->constraints.clear();
    for (const LinearInequalityConstraint& c : problem.inequality_constraints) qp->constraints.push_back(c.ShiftTo(variables));
    return e;
  }

  // Batched entry point: problems of identical shape (dimension, equality rows, inequality count), one start point
  // each. The QPs of all active problems are solved in one device call per outer iteration.
  static std::vector<NLSSolverOutputs> SolveBatch(const std::vector<const Problem*>& problems, const Params& params,
                                                  const std::vector<Vector>& initial_variables,
                                                  const Retraction& retraction = nullptr, int device = 0,
                                                  const BatchExitCallback& exit_callback = nullptr) {
    CheckParams(params);
    const std::size_t B = problems.size();
    if (initial_variables.size() != B) throw std::invalid_argument("one start point per problem");
    struct State {
      Vector x, cand, dx;
      double lambda, penalty;
      OptimizerState state = OptimizerState::NOMINAL;
      bool done = false;
      QP qp;
      Errors errors_pre;
    };
    std::vector<State> st(B);
    std::vector<NLSSolverOutputs> out(B);
    for (std::size_t i = 0; i < B; ++i) {
      if (!problems[i]) throw std::invalid_argument("Must have a valid problem");
      if (initial_variables[i].rows() != problems[i]->dimension) throw std::invalid_argument("variables / dimension mismatch");
      st[i].x = initial_variables[i];
      st[i].lambda = params.lambda_initial;
      st[i].penalty = params.equality_penalty_initial;
    }
    QPInteriorPointSolver::Params qp_params;  // nonlinear.cc:226-239
    qp_params.max_iterations = params.max_qp_iterations;
    qp_params.termination_kkt_tol = params.termination_kkt_tolerance;
    qp_params.initial_mu = 1.0;
    qp_params.sigma = 0.1;
    qp_params.initialize_mu_with_complementarity = false;
    for (int iter = 0; iter < params.max_iterations; ++iter) {
      std::vector<std::size_t> active;
      std::vector<QP> qps;
      for (std::size_t i = 0; i < B; ++i) {
        if (st[i].done) continue;
        st[i].errors_pre = LinearizeAndFillQP(st[i].x, st[i].lambda, *problems[i], &st[i].qp);
        active.push_back(i);
        qps.push_back(st[i].qp);
      }
      if (active.empty()) break;
      const Problem& first = *problems[active[0]];
      // Solver choice of the reference (nonlinear.cc:83-89): equality constraints and no inequalities -> QPNullSpaceSolver
      const bool null_space = first.inequality_constraints.empty() && !first.equality_constraints.empty();
      qp_params.initial_guess_method = first.equality_constraints.empty() ? InitialGuessMethod::NAIVE
                                                                          : InitialGuessMethod::SOLVE_EQUALITY_CONSTRAINED;
      std::vector<QPInteriorPointSolverOutputs> sol;
      std::vector<int> ns_status;
      std::vector<Vector> ns_x;
      if (null_space) ns_status = QPNullSpaceSolver::SolveBatch(qps, &ns_x, device);
      else sol = QPInteriorPointSolver::SolveBatch(qps, qp_params, device, nullptr, params.record_qp_iterations);
      for (std::size_t a = 0; a < active.size(); ++a) {
        const std::size_t i = active[a];
        State& s = st[i];
        const Problem& prob = *problems[i];
        const int n = prob.dimension;
        QPInteriorPointSolverOutputs q_ns;  // null-space: one "iteration" per solve (NumQPIterations, structs.cc:208-219)
        q_ns.num_iterations = 1;
        q_ns.termination_state = QPInteriorPointTerminationState::SATISFIED_KKT_TOL;
        const QPInteriorPointSolverOutputs& q = null_space ? q_ns : sol[a];
        if (null_space) {
          if (ns_status[a] == MO_QP_NULLSPACE_NOT_POSITIVE_DEFINITE) {  // QPWasIndefinite, nonlinear.cc:103-105
            out[i].termination_state = NLSTerminationState::QP_INDEFINITE;
            s.done = true;
            continue;
          }
          if (ns_status[a] != MO_QP_NULLSPACE_SUCCESS) {
            out[i].termination_state = NLSTerminationState::QP_FAILED;
            s.done = true;
            continue;
          }
          s.dx = ns_x[a];
        } else {
          // per-problem numerical trouble never aborts the batch: 2 = FailedFactorization (qp.cc:303-307), 3 = F_ASSERT(s > 0) (:285)
          if (q.batch_status == MO_QP_FAILED_FACTORIZATION || q.batch_status == MO_QP_NONPOSITIVE_SLACK) {
            out[i].termination_state = NLSTerminationState::QP_FAILED;
            out[i].qp_failure_status = q.batch_status;
            s.done = true;
            continue;
          }
          s.dx.assign(q.variables.begin(), q.variables.begin() + n);
        }
        // ComputeQPCostDerivative, nonlinear.cc:457-482
        DirectionalDerivatives dd{};
        for (int j = 0; j < n; ++j) dd.d_f += s.qp.c[j] * s.dx[j];
        for (int r = 0; r < s.qp.A_eq.rows(); ++r) {
          const double b = s.qp.b_eq[r], sign = b > 0 ? 1.0 : (b < 0 ? -1.0 : 0.0);
          double dot = 0;
          for (int j = 0; j < n; ++j) dot += s.qp.A_eq(r, j) * s.dx[j];
          dd.d_equality += sign * dot;
        }
        // SelectPenalty, nonlinear.cc:112-119, 484-502: |y|_inf from the interior point solve (18.32), or, with the
        // null-space solver (no multipliers), inequality (18.36)
        if (!prob.equality_constraints.empty()) {
          std::optional<double> new_penalty;
          if (!null_space && q.lagrange_multipliers) {
            new_penalty = q.lagrange_multipliers->l_infinity;
          } else if (null_space) {
            double l1 = 0, cdx = 0, quad = 0;
            for (int r = 0; r < s.qp.b_eq.rows(); ++r) l1 += std::abs(s.qp.b_eq[r]);
            l1 = std::max(l1, std::numeric_limits<double>::epsilon());
            for (int j = 0; j < n; ++j) cdx += s.qp.c[j] * s.dx[j];
            for (int r = 0; r < n; ++r) {
              double acc = 0;
              for (int j = 0; j < n; ++j) acc += (r >= j ? s.qp.G(r, j) : s.qp.G(j, r)) * s.dx[j];  // selfadjointView<Lower>
              quad += s.dx[r] * acc;
            }
            new_penalty = (cdx + 0.5 * std::max(0.0, quad)) / ((1 - params.equality_penalty_rho) * l1);
          }
          if (new_penalty && *new_penalty > s.penalty) s.penalty = *new_penalty * params.equality_penalty_scale_factor;
        }
        std::vector<LineSearchStep> steps;
        const StepSizeSelectionResult step = SelectStepSize(prob, params, s.x, s.dx, s.errors_pre, dd, s.penalty, retraction,
                                                            &s.cand, &steps);
        const double old_lambda = s.lambda;
        const OptimizerState old_state = s.state;
        const std::optional<NLSTerminationState> maybe_exit =
            UpdateLambdaAndCheckExitConditions(params, step, s.errors_pre, s.penalty, steps, &s.lambda, &s.state, &s.x, &s.cand);
        out[i].iterations.push_back(NLSIteration{iter, old_state, old_lambda, s.errors_pre, q.termination_state,
                                                 q.num_iterations, dd, s.penalty, step, std::move(steps), {}});
        if (null_space) out[i].iterations.back().qp_outputs = QPNullSpaceTerminationState::SUCCESS;
        else out[i].iterations.back().qp_outputs = q;
        if (exit_callback && !exit_callback(i, out[i].iterations.back()) && !maybe_exit) {  // nonlinear.cc:145-151
          out[i].termination_state = NLSTerminationState::USER_CALLBACK;
          s.done = true;
        }
        if (maybe_exit) {
          out[i].termination_state = *maybe_exit;
          s.done = true;
        }
      }
    }
    for (std::size_t i = 0; i < B; ++i) out[i].variables = st[i].x;  // not done: MAX_ITERATIONS (the default)
    return out;
  }

 private:
  static void CheckParams(const Params& p) {  // nonlinear.cc:49-73
    // the same condition list as check_nls_params of the C ABI (capi.cu), i.e. the reference's F_ASSERTs
    const bool ok =
        p.max_iterations >= 0 && p.max_qp_iterations >= 1 && p.termination_kkt_tolerance > 0 && p.absolute_exit_tol > 0 &&
        p.max_line_search_iterations >= 0 && p.relative_exit_tol >= 0 && p.relative_exit_tol <= 1 &&
        p.absolute_first_derivative_tol >= 0 && p.armijo_search_tau > 0 && p.armijo_search_tau < 1 &&
        p.equality_penalty_initial >= 0 && p.equality_penalty_scale_factor >= 1.0 && p.equality_penalty_rho >= 0 &&
        p.equality_penalty_rho < 1 && p.max_lambda >= 0 && p.min_lambda <= p.max_lambda && p.lambda_initial >= p.min_lambda &&
        p.lambda_initial <= p.max_lambda && p.lambda_failure_init >= 0 && p.lambda_decrease_on_success >= 0 &&
        p.lambda_decrease_on_success < 1.0 && p.lambda_decrease_on_restore >= 0 && p.lambda_decrease_on_restore < 1.0;
    if (!ok) throw std::invalid_argument("invalid ConstrainedNonlinearLeastSquares::Params (nonlinear.cc:49-73)");
  }

  // nonlinear.cc:531-538
  static std::optional<double> QuadraticApproxMinimum(double phi_0, double phi_prime_0, double alpha_0, double phi_alpha_0) {
    const double numerator = phi_alpha_0 - phi_prime_0 * alpha_0 - phi_0;
    if (phi_prime_0 > 0 || numerator <= 0) return std::nullopt;
    return -phi_prime_0 * alpha_0 * alpha_0 / (2.0 * numerator);
  }
  // nonlinear.cc:563-578 (explicit inverse of the 2x2 system) and :594-605
  static std::optional<double> CubicApproxMinimum(double phi_0, double phi_prime_0, double alpha_0, double phi_alpha_0,
                                                  double alpha_1, double phi_alpha_1) {
    const double a00 = alpha_0 * alpha_0 * alpha_0, a01 = alpha_0 * alpha_0, a10 = alpha_1 * alpha_1 * alpha_1,
                 a11 = alpha_1 * alpha_1;
    const double r0 = phi_alpha_0 - phi_0 - phi_prime_0 * alpha_0, r1 = phi_alpha_1 - phi_0 - phi_prime_0 * alpha_1;
    const double det = a00 * a11 - a01 * a10;
    const double a = (a11 * r0 - a01 * r1) / det, b = (-a10 * r0 + a00 * r1) / det;
    const double arg_sqrt = b * b - 3 * a * phi_prime_0;
    if (a == 0.0 || arg_sqrt < -1.0e-12) return std::nullopt;
    return (-b + std::sqrt(std::max(arg_sqrt, 0.))) / (3 * a);
  }

  // nonlinear.cc:351-416 (armijo_c1 = 1e-4, :121)
  static StepSizeSelectionResult SelectStepSize(const Problem& prob, const Params& params, const Vector& x, const Vector& dx,
                                                const Errors& errors_pre, const DirectionalDerivatives& derivatives,
                                                double penalty, const Retraction& retraction, Vector* cand,
                                                std::vector<LineSearchStep>* steps) {
    constexpr double armijo_c1 = 1.0e-4;
    const double directional_derivative = derivatives.Total(penalty);
    double alpha = 1;
    for (int iter = 0; iter <= params.max_line_search_iterations; ++iter) {
      if (params.line_search_strategy == LineSearchStrategy::POLYNOMIAL_APPROXIMATION) {
        if (iter > 0) {
          std::optional<double> new_alpha;
          if (iter == 1) {
            new_alpha = QuadraticApproxMinimum(errors_pre.Total(penalty), directional_derivative, steps->back().alpha,
                                               steps->back().errors.Total(penalty));
          } else {
            const LineSearchStep& s2 = (*steps)[steps->size() - 2];
            const LineSearchStep& s1 = steps->back();
            new_alpha = CubicApproxMinimum(errors_pre.Total(penalty), directional_derivative, s2.alpha,
                                           s2.errors.Total(penalty), s1.alpha, s1.errors.Total(penalty));
          }
          if (!new_alpha || !std::isfinite(*new_alpha) || *new_alpha <= 0.0 || *new_alpha >= alpha)
            return StepSizeSelectionResult::FAILURE_INVALID_ALPHA;
          alpha = *new_alpha;
        }
      } else if (iter > 0) {
        alpha = alpha * params.armijo_search_tau;
      }
      *cand = x;  // RetractCandidateVars, nonlinear.cc:160-168
      if (retraction) retraction(*cand, dx, alpha);
      else for (int j = 0; j < cand->rows(); ++j) (*cand)[j] += dx[j] * alpha;
      const Errors errors_step = EvaluateNonlinearErrors(prob, *cand);
      steps->push_back(LineSearchStep{alpha, errors_step});
      if (errors_step.ContainsInvalidValues()) return StepSizeSelectionResult::FAILURE_NON_FINITE_COST;
      if (derivatives.LInfinity() < params.absolute_first_derivative_tol) return StepSizeSelectionResult::FIRST_ORDER_SATISFIED;
      if (directional_derivative > 0) return StepSizeSelectionResult::POSITIVE_DERIVATIVE;
      if (errors_step.Total(penalty) <= errors_pre.Total(penalty) + directional_derivative * alpha * armijo_c1)
        return StepSizeSelectionResult::SUCCESS;
    }
    return StepSizeSelectionResult::MAX_ITERATIONS;
  }

  // nonlinear.cc:296-343
  static std::optional<NLSTerminationState> UpdateLambdaAndCheckExitConditions(
      const Params& params, StepSizeSelectionResult step_result, const Errors& initial_errors, double penalty,
      const std::vector<LineSearchStep>& steps, double* lambda, OptimizerState* state, Vector* x, Vector* cand) {
    if (step_result == StepSizeSelectionResult::SUCCESS) {
      x->swap(*cand);
      if (*state == OptimizerState::ATTEMPTING_RESTORE_LM) *lambda = std::max(*lambda * params.lambda_decrease_on_restore, params.min_lambda);
      else *lambda = std::max(*lambda * params.lambda_decrease_on_success, params.min_lambda);
      *state = OptimizerState::NOMINAL;
      const LineSearchStep& final_step = steps.back();
      if (final_step.errors.LInfinity() < params.absolute_exit_tol) return NLSTerminationState::SATISFIED_ABSOLUTE_TOL;
      if (final_step.errors.Total(penalty) > initial_errors.Total(penalty) * (1 - params.relative_exit_tol))
        return NLSTerminationState::SATISFIED_RELATIVE_TOL;
    } else if (step_result == StepSizeSelectionResult::FIRST_ORDER_SATISFIED) {
      return NLSTerminationState::SATISFIED_FIRST_ORDER_TOL;
    } else if (step_result == StepSizeSelectionResult::MAX_ITERATIONS ||
               step_result == StepSizeSelectionResult::POSITIVE_DERIVATIVE) {
      if (*state == OptimizerState::NOMINAL) {
        *lambda = std::max(params.lambda_failure_init, *lambda * 10.0);
        *state = OptimizerState::ATTEMPTING_RESTORE_LM;
      } else {
        *lambda *= 10.0;
      }
      if (*lambda > params.max_lambda) return NLSTerminationState::MAX_LAMBDA;
    }
    return std::nullopt;
  }

  const Problem* p_;
  Retraction custom_retraction_;
  ExitCallback user_exit_callback_;
  Vector variables_;
};

}  // namespace mini_opt_b200
