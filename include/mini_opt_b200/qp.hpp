// C++ host API of mini_opt_b200: the reference's class surface for the QP hot path (gareth-cross/mini_opt,
// include/mini_opt/qp.hpp:28-205, include/mini_opt/structs.hpp:24-134) over the C ABI of mini_opt_b200.h, plus the
// batched entry point. Header only; link with -lminiopt_b200. No Eigen dependency: `Matrix` / `Vector` are minimal
// column-major containers with Eigen::MatrixXd's storage order and accessors, so swapping the two aliases for
// Eigen::MatrixXd / Eigen::VectorXd gives the drop-in described in INTEGRATION.md.
#pragma once

#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <limits>
#include <optional>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "mini_opt_b200.h"

namespace mini_opt_b200 {

struct Vector : std::vector<double> {
  using std::vector<double>::vector;
  int rows() const { return static_cast<int>(size()); }
};

struct Matrix {  // column-major like Eigen::MatrixXd
  Matrix() = default;
  Matrix(int rows, int cols) : rows_(rows), cols_(cols), v_(static_cast<std::size_t>(rows) * cols, 0.0) {}
  void resize(int rows, int cols) { rows_ = rows; cols_ = cols; v_.assign(static_cast<std::size_t>(rows) * cols, 0.0); }
  double& operator()(int r, int c) { return v_[static_cast<std::size_t>(c) * rows_ + r]; }
  double operator()(int r, int c) const { return v_[static_cast<std::size_t>(c) * rows_ + r]; }
  int rows() const { return rows_; }
  int cols() const { return cols_; }
  const double* data() const { return v_.data(); }

 private:
  int rows_ = 0, cols_ = 0;
  std::vector<double> v_;
};

// qp.hpp:28-70: a * x[variable] + b >= 0
struct LinearInequalityConstraint {
  int variable;
  double a, b;
  LinearInequalityConstraint(int variable_, double a_, double b_) : variable(variable_), a(a_), b(b_) {}
  bool IsFeasible(double x) const { return a * x + b >= 0.0; }                        // qp.hpp:38-41
  double ClampX(double x) const { return a < 0 ? (x < b / -a ? x : b / -a) : (x > -b / a ? x : -b / a); }  // qp.hpp:43-53
  LinearInequalityConstraint ShiftTo(const Vector& x) const { return {variable, a, a * x[variable] + b}; }  // qp.hpp:57-59
};

// qp.hpp:77-92: Var(i) <= u, Var(i) >= l
struct Var {
  explicit Var(int variable_) : variable(variable_) {}
  LinearInequalityConstraint operator<=(double value) const { return {variable, -1.0, value}; }
  LinearInequalityConstraint operator>=(double value) const { return {variable, 1.0, -value}; }
  int variable;
};

// qp.hpp:104-124: minimise 1/2 x'Gx + c'x  s.t.  A_eq x + b_eq = 0, a_i x[v_i] + b_i >= 0
struct QP {
  QP() = default;
  explicit QP(int n) : G(n, n), c(static_cast<std::size_t>(n), 0.0) {}
  Matrix G;
  Vector c;
  Matrix A_eq;
  Vector b_eq;
  std::vector<LinearInequalityConstraint> constraints;
};

enum class BarrierStrategy { COMPLEMENTARITY = 0, FIXED_DECREASE = 1, PREDICTOR_CORRECTOR = 2 };        // structs.hpp:24-31
enum class InitialGuessMethod { NAIVE = 0, SOLVE_EQUALITY_CONSTRAINED = 1, USER_PROVIDED = 2 };          // structs.hpp:34-41
enum class QPInteriorPointTerminationState { SATISFIED_KKT_TOL = 0, MAX_ITERATIONS = 1 };                // structs.hpp:97-102

// ---- number formatting of the reference's logs ----
// The reference prints through {fmt} ("{}" = shortest round-trip digits, fixed notation for 1e-4 <= |v| < 1e16, else
// d.ddde+XX) and serializes through nlohmann::json (same digits, fixed up to 1e15 with a trailing ".0" on integers,
// NaN / Inf as null). Both are reproduced on top of std::to_chars so that ToString() / ToJson() of the batched results
// are byte-identical to the reference's on equal values (tests/golden/logging_reference.txt).
namespace detail {
inline std::string FormatShortest(double v, bool json) {
  if (std::isnan(v)) return json ? "null" : (std::signbit(v) ? "-nan" : "nan");
  if (std::isinf(v)) return json ? "null" : (v < 0 ? "-inf" : "inf");
  if (v == 0) return std::signbit(v) ? (json ? "-0.0" : "-0") : (json ? "0.0" : "0");
  char buf[40];
  const auto res = std::to_chars(buf, buf + sizeof(buf), std::fabs(v), std::chars_format::scientific);
  std::string digits;
  int exp10 = 0;
  {
    const std::string t(buf, res.ptr);
    const std::size_t e = t.find('e');
    for (std::size_t i = 0; i < e; ++i)
      if (t[i] != '.') digits += t[i];
    exp10 = std::stoi(t.substr(e + 1));
  }
  const int nd = static_cast<int>(digits.size()), point = exp10 + 1;  // value = 0.d1d2... x 10^point
  std::string out = v < 0 ? "-" : "";
  const int fixed_max = json ? 15 : 16;
  if (point > -4 && point <= fixed_max) {
    if (point <= 0) {
      out += "0." + std::string(static_cast<std::size_t>(-point), '0') + digits;
    } else if (nd <= point) {
      out += digits + std::string(static_cast<std::size_t>(point - nd), '0');
      if (json) out += ".0";
    } else {
      out += digits.substr(0, static_cast<std::size_t>(point)) + "." + digits.substr(static_cast<std::size_t>(point));
    }
  } else {
    out += digits.substr(0, 1);
    if (nd > 1) out += "." + digits.substr(1);
    char eb[16];
    std::snprintf(eb, sizeof(eb), "e%c%02d", exp10 < 0 ? '-' : '+', exp10 < 0 ? -exp10 : exp10);
    out += eb;
  }
  return out;
}
inline std::string Fmt(double v) { return FormatShortest(v, false); }
inline std::string Json(double v) { return FormatShortest(v, true); }
inline std::string Printf(const char* format, double v) {
  char buf[400];
  std::snprintf(buf, sizeof(buf), format, v);
  return buf;
}
}  // namespace detail

struct KKTError {  // structs.hpp:68-78
  double r_dual = 0, r_comp = 0, r_primal_eq = 0, r_primal_ineq = 0;
  double Max() const {
    const double a = r_dual > r_comp ? r_dual : r_comp, b = r_primal_eq > r_primal_ineq ? r_primal_eq : r_primal_ineq;
    return a > b ? a : b;
  }
};
struct QPLagrangeMultipliers { double min, l_infinity; };  // structs.hpp:105-113
struct AlphaValues { double primal{1.0}, dual{1.0}; };  // structs.hpp:44-49
struct IPIterationOutputs {                               // structs.hpp:52-64
  double mu{0.0};
  AlphaValues alpha{};
  AlphaValues alpha_probe{std::numeric_limits<double>::quiet_NaN(), std::numeric_limits<double>::quiet_NaN()};
  double mu_affine{std::numeric_limits<double>::quiet_NaN()};
};
// One Newton step of one problem (structs.hpp:81-94) = one row of mo_qp_out.iter_log.
struct QPInteriorPointIteration {
  KKTError kkt_initial{}, kkt_final{};
  IPIterationOutputs ip_outputs{};
  static QPInteriorPointIteration FromLogRow(const double* row) {  // MO_QP_ITER_LOG_STRIDE doubles
    QPInteriorPointIteration it;
    it.kkt_initial = KKTError{row[0], row[1], row[2], row[3]};
    it.kkt_final = KKTError{row[4], row[5], row[6], row[7]};
    it.ip_outputs.mu = row[8];
    it.ip_outputs.alpha = AlphaValues{row[9], row[10]};
    it.ip_outputs.alpha_probe = AlphaValues{row[11], row[12]};
    it.ip_outputs.mu_affine = row[13];
    return it;
  }
  std::string ToString() const {  // structs.cc:122-148
    using detail::Fmt;
    std::string r = "Iteration summary: ||kkt|| max: " + Fmt(kkt_initial.Max()) + " --> " + Fmt(kkt_final.Max()) +
                    ", mu = " + Fmt(ip_outputs.mu) + ", a_p = " + Fmt(ip_outputs.alpha.primal) +
                    ", a_d = " + Fmt(ip_outputs.alpha.dual) + "\n";
    if (!std::isnan(ip_outputs.mu_affine))
      r += " Probe alphas: a_p = " + Fmt(ip_outputs.alpha_probe.primal) + ", a_d = " + Fmt(ip_outputs.alpha_probe.dual) +
           ", mu_affine = " + Fmt(ip_outputs.mu_affine) + "\n";
    r += " KKT errors, L2:\n r_dual = " + Fmt(kkt_initial.r_dual) + " --> " + Fmt(kkt_final.r_dual) +
         "\n r_comp = " + Fmt(kkt_initial.r_comp) + " --> " + Fmt(kkt_final.r_comp) +
         "\n r_p_eq = " + Fmt(kkt_initial.r_primal_eq) + " --> " + Fmt(kkt_final.r_primal_eq) +
         "\n r_p_ineq = " + Fmt(kkt_initial.r_primal_ineq) + " --> " + Fmt(kkt_final.r_primal_ineq) + "\n";
    return r;
  }
  std::string ToJson() const {  // serialization.cc:71-74 (keys in nlohmann's alphabetical order)
    using detail::Json;
    auto kkt = [](const KKTError& e) {
      return "{\"r_comp\":" + Json(e.r_comp) + ",\"r_dual\":" + Json(e.r_dual) + ",\"r_primal_eq\":" + Json(e.r_primal_eq) +
             ",\"r_primal_ineq\":" + Json(e.r_primal_ineq) + "}";
    };
    auto alpha = [](const AlphaValues& a) { return "{\"dual\":" + Json(a.dual) + ",\"primal\":" + Json(a.primal) + "}"; };
    return "{\"ip_outputs\":{\"alpha\":" + alpha(ip_outputs.alpha) + ",\"alpha_probe\":" + alpha(ip_outputs.alpha_probe) +
           ",\"mu\":" + Json(ip_outputs.mu) + ",\"mu_affine\":" + Json(ip_outputs.mu_affine) + "},\"kkt_final\":" + kkt(kkt_final) +
           ",\"kkt_initial\":" + kkt(kkt_initial) + "}";
  }
};
inline const char* ToString(QPInteriorPointTerminationState s) {  // structs.cc:26-36
  return s == QPInteriorPointTerminationState::SATISFIED_KKT_TOL ? "SATISFIED_KKT_TOL" : "MAX_ITERATIONS";
}
inline std::ostream& operator<<(std::ostream& os, QPInteriorPointTerminationState s) { return os << ToString(s); }

struct FailedFactorization : std::runtime_error { using std::runtime_error::runtime_error; };  // qp.hpp:331-333

// Per-problem result of the batched solve. `variables` = [x | s | y | z] (qp.cc:548-582).
struct QPInteriorPointSolverOutputs {
  QPInteriorPointTerminationState termination_state = QPInteriorPointTerminationState::MAX_ITERATIONS;
  int num_iterations = 0;      // == iterations.size() of the reference's outputs
  // the reference's per-step log (structs.hpp:120-122), filled from mo_qp_out.iter_log when the solve was asked to record
  // it (Solve: always, like the reference; SolveBatch: `record_iterations`, 112 bytes per step come back from the device)
  std::vector<QPInteriorPointIteration> iterations;
  KKTError kkt_final;          // iterations.back().kkt_final
  std::optional<QPLagrangeMultipliers> lagrange_multipliers;
  int batch_status = MO_QP_MAX_ITERATIONS;  // MO_QP_*: 2 / 3 where the reference would have thrown
  Vector variables;
  int n = 0, k = 0, m = 0;
  double x(int i) const { return variables[i]; }
  double s(int i) const { return variables[n + i]; }
  double y(int i) const { return variables[n + m + i]; }
  double z(int i) const { return variables[n + m + k + i]; }
  std::string ToJson() const {  // serialization.cc:76-77
    std::string r = "{\"iterations\":[";
    for (std::size_t i = 0; i < iterations.size(); ++i) r += (i ? "," : "") + iterations[i].ToJson();
    r += "],\"lagrange_multipliers\":";
    r += lagrange_multipliers ? "{\"l_infinity\":" + detail::Json(lagrange_multipliers->l_infinity) + ",\"min\":" +
                                    detail::Json(lagrange_multipliers->min) + "}"
                              : std::string("null");
    return r + ",\"termination_state\":\"" + ToString(termination_state) + "\"}";
  }
};

class QPInteriorPointSolver {
 public:
  struct Params {  // qp.hpp:134-164, same defaults
    double initial_mu{1.0};
    double sigma{0.5};
    double termination_kkt_tol{1.0e-9};
    double termination_complementarity_tol{1.0e-6};
    int max_iterations{10};
    BarrierStrategy barrier_strategy{BarrierStrategy::COMPLEMENTARITY};
    bool decrease_mu_only_on_small_error{false};
    InitialGuessMethod initial_guess_method{InitialGuessMethod::NAIVE};
    bool initialize_mu_with_complementarity{false};
  };

  QPInteriorPointSolver() = default;
  explicit QPInteriorPointSolver(const QP* problem) { Setup(problem); }

  // qp.hpp:173-174: the problem is not copied and must outlive the solver.
  void Setup(const QP* problem) {
    if (!problem) throw std::invalid_argument("Must pass a non-null problem");
    Check(*problem);
    p_ = problem;
  }

  // One problem = a batch of one (host buffers; H2D, solve, D2H inside). Throws FailedFactorization like qp.cc:303-307.
  QPInteriorPointSolverOutputs Solve(const Params& params, int device = 0) {
    if (!p_) throw std::invalid_argument("Must have a valid problem");
    std::vector<Vector> init;
    if (params.initial_guess_method == InitialGuessMethod::USER_PROVIDED) init.push_back(variables_);
    std::vector<QPInteriorPointSolverOutputs> out = SolveBatch({*p_}, params, device, init.empty() ? nullptr : &init, true);
    if (out[0].batch_status == MO_QP_FAILED_FACTORIZATION) throw FailedFactorization("Failed to solve self-adjoint system");
    if (out[0].batch_status == MO_QP_NONPOSITIVE_SLACK) throw std::runtime_error("Some slack variables s <= 0");
    variables_ = out[0].variables;
    return out[0];
  }

  const Vector& variables() const { return variables_; }
  const QP& problem() const { return *p_; }
  // qp.cc:221: the start point [x | s | y | z] for InitialGuessMethod::USER_PROVIDED (n + 2m + k entries)
  void SetVariables(const Vector& v) {
    if (!p_) throw std::invalid_argument("Must have a valid problem");
    const std::size_t want = p_->G.rows() + 2 * p_->constraints.size() + p_->A_eq.rows();
    if (v.size() != want) throw std::invalid_argument("SetVariables: wrong size");
    variables_ = v;
  }
  // x_block() / s_block() / y_block() / z_block() of the reference (qp.hpp:187-198) as copies of the slices
  Vector x_block() const { return Slice(0, p_->G.rows()); }
  Vector s_block() const { return Slice(p_->G.rows(), static_cast<int>(p_->constraints.size())); }
  Vector y_block() const { return Slice(p_->G.rows() + static_cast<int>(p_->constraints.size()), p_->A_eq.rows()); }
  Vector z_block() const {
    return Slice(p_->G.rows() + static_cast<int>(p_->constraints.size()) + p_->A_eq.rows(), static_cast<int>(p_->constraints.size()));
  }

  // Batched entry point: every problem must have the same (n, k, m). Per-problem failures are reported in
  // `batch_status`, never thrown. `flags_box_pairs`: the constraints of every problem are the interleaved pairs
  // Var(i) >= lo_i, Var(i) <= hi_i for i = 0..n-1 (selects the tuned kernels).
  // `initial_variables`: one [x | s | y | z] start point per problem, required for InitialGuessMethod::USER_PROVIDED
  // (the batched SetVariables) and ignored otherwise.
  static std::vector<QPInteriorPointSolverOutputs> SolveBatch(const std::vector<QP>& qps, const Params& params,
                                                              int device = 0,
                                                              const std::vector<Vector>* initial_variables = nullptr,
                                                              bool record_iterations = false) {
    std::vector<QPInteriorPointSolverOutputs> result(qps.size());
    if (qps.empty()) return result;
    const bool user_guess = params.initial_guess_method == InitialGuessMethod::USER_PROVIDED;
    if (user_guess && (!initial_variables || initial_variables->size() != qps.size()))
      throw std::invalid_argument("USER_PROVIDED needs SetVariables / one initial [x|s|y|z] per problem");
    const int n = qps[0].G.rows(), k = qps[0].A_eq.rows(), m = static_cast<int>(qps[0].constraints.size());
    const std::int64_t B = static_cast<std::int64_t>(qps.size()), stride = mo_qp_vars_stride(n, k, m);
    std::vector<double> G(B * n * n), c(B * n), A(B * k * n + 1), b(B * k + 1), ia(B * m + 1), ib(B * m + 1);
    std::vector<double> vars(B * stride), kkt(B * 4), lag(B * 2);
    std::vector<std::int32_t> iv(B * m + 1), status(B), iters(B);
    const std::int64_t log_stride = static_cast<std::int64_t>(params.max_iterations > 0 ? params.max_iterations : 0) * MO_QP_ITER_LOG_STRIDE;
    std::vector<double> log(record_iterations ? B * log_stride : 0);
    bool box = m == 2 * n;
    for (std::int64_t p = 0; p < B; ++p) {
      const QP& q = qps[p];
      Check(q);
      if (q.G.rows() != n || q.A_eq.rows() != k || static_cast<int>(q.constraints.size()) != m)
        throw std::invalid_argument("SolveBatch needs problems of identical shape");
      if (user_guess) {
        const Vector& v0 = (*initial_variables)[p];
        if (static_cast<std::int64_t>(v0.size()) != stride) throw std::invalid_argument("initial variables: wrong size");
        for (std::int64_t i = 0; i < stride; ++i) vars[p * stride + i] = v0[i];
      }
      for (int i = 0; i < n * n; ++i) G[p * n * n + i] = q.G.data()[i];
      for (int i = 0; i < n; ++i) c[p * n + i] = q.c[i];
      for (int i = 0; i < k * n; ++i) A[p * k * n + i] = q.A_eq.data()[i];
      for (int i = 0; i < k; ++i) b[p * k + i] = q.b_eq[i];
      for (int i = 0; i < m; ++i) {
        const LinearInequalityConstraint& con = q.constraints[i];
        iv[p * m + i] = con.variable; ia[p * m + i] = con.a; ib[p * m + i] = con.b;
        box = box && con.variable == i / 2 && con.a == ((i % 2 == 0) ? 1.0 : -1.0);
      }
    }
    mo_qp_batch batch{n, k, m, box ? MO_QP_FLAG_BOX_PAIRS : 0, B, G.data(), c.data(), k ? A.data() : nullptr,
                      k ? b.data() : nullptr, m ? iv.data() : nullptr, m ? ia.data() : nullptr, m ? ib.data() : nullptr,
                      nullptr};
    mo_qp_params prm;
    mo_qp_params_default(&prm);
    prm.initial_mu = params.initial_mu; prm.sigma = params.sigma;
    prm.termination_kkt_tol = params.termination_kkt_tol;
    prm.termination_complementarity_tol = params.termination_complementarity_tol;
    prm.max_iterations = params.max_iterations; prm.barrier_strategy = static_cast<int>(params.barrier_strategy);
    prm.decrease_mu_only_on_small_error = params.decrease_mu_only_on_small_error;
    prm.initial_guess_method = static_cast<int>(params.initial_guess_method);
    prm.initialize_mu_with_complementarity = params.initialize_mu_with_complementarity;
    mo_qp_out out{vars.data(), status.data(), iters.data(), kkt.data(), lag.data(), log.empty() ? nullptr : log.data(), nullptr};
    const int rc = mo_qp_ip_solve_batch_host(&batch, &prm, &out, device);
    if (rc == MO_ERR_INVALID_PARAMS) throw std::invalid_argument("invalid QPInteriorPointSolver::Params (qp.cc:76-82)");
    if (rc == MO_ERR_INVALID_ARGUMENT) throw std::invalid_argument("invalid problem");
    if (rc != MO_OK) throw std::runtime_error(std::string("mini_opt_b200 error ") + std::to_string(rc) + ": " + mo_last_error_string());
    for (std::int64_t p = 0; p < B; ++p) {
      QPInteriorPointSolverOutputs& o = result[p];
      o.batch_status = status[p];
      o.termination_state = status[p] == MO_QP_SATISFIED_KKT_TOL ? QPInteriorPointTerminationState::SATISFIED_KKT_TOL
                                                                 : QPInteriorPointTerminationState::MAX_ITERATIONS;
      o.num_iterations = iters[p];
      if (!log.empty())
        for (int it = 0; it < iters[p]; ++it)
          o.iterations.push_back(QPInteriorPointIteration::FromLogRow(log.data() + p * log_stride + it * MO_QP_ITER_LOG_STRIDE));
      o.kkt_final = KKTError{kkt[4 * p], kkt[4 * p + 1], kkt[4 * p + 2], kkt[4 * p + 3]};
      if (k > 0) o.lagrange_multipliers = QPLagrangeMultipliers{lag[2 * p], lag[2 * p + 1]};
      o.variables.assign(vars.begin() + p * stride, vars.begin() + (p + 1) * stride);
      o.n = n; o.k = k; o.m = m;
    }
    return result;
  }

 private:
  static void Check(const QP& q) {  // Setup's assertions, qp.cc:20-44
    if (q.G.rows() != q.G.cols() || q.G.rows() != q.c.rows() || q.G.rows() <= 0) throw std::invalid_argument("G must be square and match c");
    if (q.A_eq.rows() != q.b_eq.rows() || (q.A_eq.rows() > 0 && q.A_eq.cols() != q.G.rows())) throw std::invalid_argument("A_eq / b_eq mismatch");
    for (const LinearInequalityConstraint& con : q.constraints)
      if (con.variable < 0 || con.variable >= q.G.rows()) throw std::invalid_argument("constraint index out of bounds");
  }
  Vector Slice(int begin, int count) const {
    if (static_cast<int>(variables_.size()) < begin + count) return Vector(static_cast<std::size_t>(count), 0.0);
    return Vector(variables_.begin() + begin, variables_.begin() + begin + count);
  }
  const QP* p_ = nullptr;
  Vector variables_;
};

// ---------------------------------------------------------------------------------------------------------
// General (non-diagonal) linear inequalities  a . x + b >= 0  (SURVEY 8(f) row 4). The reference does not have them
// (README.md:36-37 lists them as future work), so there is no reference code path to mirror; they are provided by an EXACT
// reformulation onto the diagonal form every kernel already solves: one new variable t_j per general constraint,
//     t_j - a_j . x - b_j = 0   (an equality row),      t_j >= 0   (a diagonal inequality),
// i.e. the lifted problem has n + g variables, k + g equality rows and m + g diagonal constraints, G' = blkdiag(G, 0).
// Its KKT conditions are the original problem's with z_j (the multiplier of t_j >= 0) as the multiplier of constraint j:
// G x + c - A_eq' y - sum_j a_j z_j = 0. Sigma = z / s on the t block makes the reduced Hessian positive definite from the
// first Newton step on; the equality-constrained initial guess would factor blkdiag(G, 0) without Sigma (a zero pivot in
// the unpivoted kernels), so the start point is built on the host instead: x0 = the reference's naive start (0 clamped into
// the diagonal constraints, qp.cc:462-481), t0 = max(a . x0 + b, 1), s = t0, z = 1 / t0.
// ---------------------------------------------------------------------------------------------------------
struct GeneralInequality {
  Vector a;  // n coefficients
  double b;
  double Evaluate(const Vector& x) const {
    double v = b;
    for (std::size_t i = 0; i < a.size(); ++i) v += a[i] * x[i];
    return v;
  }
};

inline QP LiftGeneralInequalities(const QP& qp, const std::vector<GeneralInequality>& general) {
  const int n = qp.G.rows(), k = qp.A_eq.rows(), g = static_cast<int>(general.size());
  QP out(n + g);
  for (int c = 0; c < n; ++c)
    for (int r = 0; r < n; ++r) out.G(r, c) = qp.G(r, c);
  for (int i = 0; i < n; ++i) out.c[i] = qp.c[i];
  out.A_eq.resize(k + g, n + g);
  out.b_eq = Vector(static_cast<std::size_t>(k + g), 0.0);
  for (int r = 0; r < k; ++r) {
    for (int c = 0; c < n; ++c) out.A_eq(r, c) = qp.A_eq(r, c);
    out.b_eq[r] = qp.b_eq[r];
  }
  out.constraints = qp.constraints;
  for (int j = 0; j < g; ++j) {
    if (static_cast<int>(general[j].a.size()) != n) throw std::invalid_argument("GeneralInequality: a must have n entries");
    for (int c = 0; c < n; ++c) out.A_eq(k + j, c) = -general[j].a[c];
    out.A_eq(k + j, n + j) = 1.0;
    out.b_eq[k + j] = -general[j].b;
    out.constraints.push_back(Var(n + j) >= 0.0);
  }
  return out;
}

// Solves a batch of QPs with general inequalities (every problem: the same n, k, m and the same NUMBER of general
// constraints). The outputs are in the ORIGINAL problem's terms with the general constraints appended to the diagonal ones:
// variables = [x (n) | s (m + g) | y (k) | z (m + g)], n / k / m set accordingly.
inline std::vector<QPInteriorPointSolverOutputs> SolveBatchWithGeneralInequalities(
    const std::vector<QP>& qps, const std::vector<std::vector<GeneralInequality>>& general,
    QPInteriorPointSolver::Params params, int device = 0) {
  if (qps.size() != general.size()) throw std::invalid_argument("one list of general inequalities per problem");
  std::vector<QP> lifted;
  std::vector<Vector> start;
  lifted.reserve(qps.size());
  for (std::size_t p = 0; p < qps.size(); ++p) {
    const QP& q = qps[p];
    const int n = q.G.rows(), k = q.A_eq.rows(), m = static_cast<int>(q.constraints.size()), g = static_cast<int>(general[p].size());
    lifted.push_back(LiftGeneralInequalities(q, general[p]));
    Vector x0(static_cast<std::size_t>(n), 0.0);
    for (const LinearInequalityConstraint& con : q.constraints) x0[con.variable] = con.ClampX(x0[con.variable]);
    Vector v(static_cast<std::size_t>(n + g + 2 * (m + g) + k + g), 0.0);  // [x, t | s | y | z] of the lifted problem
    for (int i = 0; i < n; ++i) v[i] = x0[i];
    double* sv = v.data() + n + g;
    double* zv = v.data() + n + g + (m + g) + (k + g);
    for (int i = 0; i < m; ++i) {
      const LinearInequalityConstraint& con = q.constraints[i];
      const double s = con.a * x0[con.variable] + con.b;
      sv[i] = s > 1.0e-9 ? s : 1.0e-9;
      zv[i] = 1.0 / sv[i];
    }
    for (int j = 0; j < g; ++j) {
      const double t = general[p][j].Evaluate(x0);
      v[n + j] = t > 1.0 ? t : 1.0;
      sv[m + j] = v[n + j];
      zv[m + j] = 1.0 / sv[m + j];
    }
    start.push_back(std::move(v));
  }
  params.initial_guess_method = InitialGuessMethod::USER_PROVIDED;
  std::vector<QPInteriorPointSolverOutputs> sol = QPInteriorPointSolver::SolveBatch(lifted, params, device, &start);
  for (std::size_t p = 0; p < qps.size(); ++p) {
    const QP& q = qps[p];
    const int n = q.G.rows(), k = q.A_eq.rows(), m = static_cast<int>(q.constraints.size()), g = static_cast<int>(general[p].size());
    const Vector& lv = sol[p].variables;
    Vector v;
    v.insert(v.end(), lv.begin(), lv.begin() + n);                                    // x
    v.insert(v.end(), lv.begin() + n + g, lv.begin() + n + g + m + g);                // s (diagonal, then general)
    v.insert(v.end(), lv.begin() + n + g + m + g, lv.begin() + n + g + m + g + k);    // y of the original equality rows
    v.insert(v.end(), lv.begin() + n + g + m + g + k + g, lv.end());                  // z
    sol[p].variables = std::move(v);
    sol[p].n = n; sol[p].k = k; sol[p].m = m + g;
    if (k == 0) sol[p].lagrange_multipliers.reset();
    else {  // the summary over the ORIGINAL equality rows only (qp.cc:539-546)
      double mn = sol[p].y(0), li = std::fabs(sol[p].y(0));
      for (int r = 1; r < k; ++r) { mn = sol[p].y(r) < mn ? sol[p].y(r) : mn; li = std::fabs(sol[p].y(r)) > li ? std::fabs(sol[p].y(r)) : li; }
      sol[p].lagrange_multipliers = QPLagrangeMultipliers{mn, li};
    }
  }
  return sol;
}

enum class QPNullSpaceTerminationState { SUCCESS = 0, NOT_POSITIVE_DEFINITE = 1 };  // structs.hpp:137-142
inline const char* ToString(QPNullSpaceTerminationState s) {                         // structs.cc:38-48
  return s == QPNullSpaceTerminationState::SUCCESS ? "SUCCESS" : "NOT_POSITIVE_DEFINITE";
}
inline std::ostream& operator<<(std::ostream& os, QPNullSpaceTerminationState s) { return os << ToString(s); }

// qp.hpp:304-319: equality-constrained QP by the null-space method (source/qp.cc:679-729); the inequality constraints of
// the problem are ignored like in the reference. SolveBatch: problems of identical (n, k), one device call.
class QPNullSpaceSolver {
 public:
  QPNullSpaceTerminationState Solve(const QP& p, int device = 0) {
    std::vector<Vector> x;
    const std::vector<int> st = SolveBatch({p}, &x, device);
    if (st[0] == MO_QP_NULLSPACE_RANK_DEFICIENT) throw std::runtime_error("A_eq is rank deficient");
    x_ = x[0];
    return static_cast<QPNullSpaceTerminationState>(st[0]);
  }
  const Vector& variables() const { return x_; }

  // Returns MO_QP_NULLSPACE_* per problem; x[i] = 0 unless SUCCESS.
  static std::vector<int> SolveBatch(const std::vector<QP>& qps, std::vector<Vector>* x, int device = 0) {
    std::vector<int> result(qps.size(), 0);
    x->assign(qps.size(), Vector());
    if (qps.empty()) return result;
    const int n = qps[0].G.rows(), k = qps[0].A_eq.rows();
    if (k <= 0) throw std::invalid_argument("Problem must have at least one equality constraint");
    const std::int64_t B = static_cast<std::int64_t>(qps.size());
    std::vector<double> G(B * n * n), c(B * n), A(B * k * n), b(B * k), xs(B * n);
    std::vector<std::int32_t> status(B);
    for (std::int64_t p = 0; p < B; ++p) {
      const QP& q = qps[p];
      if (q.G.rows() != n || q.G.cols() != n || q.c.rows() != n || q.A_eq.rows() != k || q.A_eq.cols() != n || q.b_eq.rows() != k)
        throw std::invalid_argument("QPNullSpaceSolver::SolveBatch needs problems of identical shape");
      for (int i = 0; i < n * n; ++i) G[p * n * n + i] = q.G.data()[i];
      for (int i = 0; i < n; ++i) c[p * n + i] = q.c[i];
      for (int i = 0; i < k * n; ++i) A[p * k * n + i] = q.A_eq.data()[i];
      for (int i = 0; i < k; ++i) b[p * k + i] = q.b_eq[i];
    }
    mo_qp_batch batch{n, k, 0, 0, B, G.data(), c.data(), A.data(), b.data(), nullptr, nullptr, nullptr, nullptr};
    const int rc = mo_qp_nullspace_solve_batch_host(&batch, xs.data(), status.data(), device);
    if (rc != MO_OK) throw std::runtime_error(std::string("mini_opt_b200 error ") + std::to_string(rc) + ": " + mo_last_error_string());
    for (std::int64_t p = 0; p < B; ++p) {
      result[p] = status[p];
      (*x)[p].assign(xs.begin() + p * n, xs.begin() + (p + 1) * n);
    }
    return result;
  }

 private:
  Vector x_;
};

}  // namespace mini_opt_b200
