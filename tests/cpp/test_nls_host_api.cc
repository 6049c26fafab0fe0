// C++ host API test for include/mini_opt_b200/nonlinear.hpp: the reference's own nonlinear test problems
// (gareth-cross/mini_opt test/nonlinear_test.cc:375-510 Rosenbrock / RosenbrockLM / inequality constrained Rosenbrock,
// :721-830 sphere with nonlinear equality constraints, test/residual_test.cc:51-182 linearization) through
// ConstrainedNonlinearLeastSquares::SolveBatch, every start point of a test as one batch.
// usage: test_nls_host_api cpu|gpu   ("cpu": no device calls -- Residual / LinearizeAndFillQP only)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>

#include "mini_opt_b200/nonlinear.hpp"

using namespace mini_opt_b200;

#define CHECK(cond)                                                                 \
  do {                                                                              \
    if (!(cond)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); std::exit(1); } \
  } while (0)

// h(x,y)'h(x,y) = (a - x)^2 + b (y - x^2)^2, a = 1, b = 100 (nonlinear_test.cc:375-387); J column-major 2x2
static std::array<double, 2> Rosenbrock(const std::array<double, 2>& p, std::array<double, 4>* J) {
  const double sb = std::sqrt(100.0);
  if (J) { (*J)[0] = -1.0; (*J)[1] = -2 * p[0] * sb; (*J)[2] = 0.0; (*J)[3] = sb; }
  return {1.0 - p[0], sb * (p[1] - p[0] * p[0])};
}
template <int N>
static std::array<double, N> SphereFunction(const std::array<double, N>& x, std::array<double, N * N>* J) {
  if (J) { J->fill(0.0); for (int i = 0; i < N; ++i) (*J)[i * N + i] = 1.0; }
  return x;
}
static std::array<double, 1> ProductExpression(const std::array<double, 2>& x, double v, std::array<double, 2>* J) {
  if (J) { (*J)[0] = x[1]; (*J)[1] = x[0]; }
  return {x[0] * x[1] - v};
}

// one line per start point in the format of oracle/ref_nls_host_golden.cc (compared by tests/test_cpp_host_api.py)
static void PrintResults(const char* name, const std::vector<NLSSolverOutputs>& out) {
  for (std::size_t i = 0; i < out.size(); ++i) {
    int qp_iters = 0;
    for (const NLSIteration& it : out[i].iterations) qp_iters += it.qp_iterations;
    std::printf("RESULT %s %zu %d %zu %d %.17g %.17g\n", name, i, static_cast<int>(out[i].termination_state),
                out[i].iterations.size(), qp_iters, out[i].variables[0], out[i].variables[1]);
  }
}

static bool SatisfiedTol(NLSTerminationState s) {  // structs.hpp:243-248
  return s == NLSTerminationState::SATISFIED_ABSOLUTE_TOL || s == NLSTerminationState::SATISFIED_RELATIVE_TOL ||
         s == NLSTerminationState::SATISFIED_FIRST_ORDER_TOL;
}

static void TestLinearization() {  // residual_test.cc:51-182, nonlinear_test.cc:109-183
  Problem problem;
  problem.dimension = 3;
  problem.costs.push_back(MakeResidual<2, 2>({2, 0}, &Rosenbrock));  // scattered, out-of-order index
  problem.equality_constraints.push_back(MakeResidual<1, 2>(
      {0, 1}, [](const std::array<double, 2>& x, std::array<double, 2>* J) { return ProductExpression(x, 4.0, J); }));
  problem.inequality_constraints.push_back(Var(1) >= -2.0);
  const Vector x = {0.5, -1.5, 2.0};
  QP qp;
  const Errors e = ConstrainedNonlinearLeastSquares::LinearizeAndFillQP(x, 0.25, problem, &qp);
  std::array<double, 4> J;
  const std::array<double, 2> r = Rosenbrock({x[2], x[0]}, &J);  // params (x2, x0)
  // J'J in global coordinates: local 0 -> global 2, local 1 -> global 0; only the lower triangle is written
  const double j00 = J[0] * J[0] + J[1] * J[1], j01 = J[0] * J[2] + J[1] * J[3], j11 = J[2] * J[2] + J[3] * J[3];
  CHECK(std::abs(qp.G(2, 2) - (j00 + 0.25)) < 1e-12 && std::abs(qp.G(0, 0) - (j11 + 0.25)) < 1e-12);
  CHECK(std::abs(qp.G(2, 0) - j01) < 1e-12 && qp.G(0, 2) == 0.0 && qp.G(1, 1) == 0.25 && qp.G(1, 0) == 0.0);
  CHECK(std::abs(qp.c[2] - (J[0] * r[0] + J[1] * r[1])) < 1e-12 && std::abs(qp.c[0] - (J[2] * r[0] + J[3] * r[1])) < 1e-12);
  CHECK(qp.c[1] == 0.0);
  CHECK(std::abs(e.f - 0.5 * (r[0] * r[0] + r[1] * r[1])) < 1e-12);
  CHECK(qp.A_eq.rows() == 1 && qp.A_eq(0, 0) == x[1] && qp.A_eq(0, 1) == x[0] && qp.A_eq(0, 2) == 0.0);
  CHECK(std::abs(qp.b_eq[0] - (x[0] * x[1] - 4.0)) < 1e-15 && std::abs(e.equality - std::abs(qp.b_eq[0])) < 1e-15);
  CHECK(qp.constraints.size() == 1 && qp.constraints[0].variable == 1 && qp.constraints[0].b == x[1] + 2.0);  // ShiftTo
  const Errors e2 = ConstrainedNonlinearLeastSquares::EvaluateNonlinearErrors(problem, x);
  CHECK(std::abs(e2.f - e.f) < 1e-15 && std::abs(e2.equality - e.equality) < 1e-15);
  bool threw = false;
  try { ConstrainedNonlinearLeastSquares::Params p; p.armijo_search_tau = 1.5; ConstrainedNonlinearLeastSquares::SolveBatch({}, p, {}); }
  catch (const std::invalid_argument&) { threw = true; }
  CHECK(threw);
}

static void TestRosenbrock(bool lm) {  // nonlinear_test.cc:391-460
  Problem problem;
  problem.costs.push_back(MakeResidual<2, 2>({0, 1}, &Rosenbrock));
  problem.dimension = 2;
  ConstrainedNonlinearLeastSquares::Params p;
  p.max_iterations = lm ? 10 : 5;
  p.max_qp_iterations = 1;
  if (lm) { p.absolute_first_derivative_tol = 1e-12; p.max_line_search_iterations = 0; }
  const std::vector<Vector> guesses = {{-5, -3}, {10, 8}, {-20, 3}, {0, -5}, {4, 0}, {100, 50}, {-35, 40}, {1000, -50}, {0.8, -0.3}};
  const std::vector<const Problem*> problems(guesses.size(), &problem);
  const std::vector<NLSSolverOutputs> out = ConstrainedNonlinearLeastSquares::SolveBatch(problems, p, guesses);
  PrintResults(lm ? "rosenbrock_lm" : "rosenbrock", out);
  for (std::size_t i = 0; i < out.size(); ++i) {
    CHECK(out[i].termination_state == NLSTerminationState::SATISFIED_ABSOLUTE_TOL);
    int qp_iters = 0;
    for (const NLSIteration& it : out[i].iterations) qp_iters += it.qp_iterations;
    CHECK(qp_iters == static_cast<int>(out[i].iterations.size()));  // NumQPIterations() == iterations.size()
    CHECK(std::abs(out[i].variables[0] - 1.0) < 1e-6 && std::abs(out[i].variables[1] - 1.0) < 1e-6);
  }
  // the single-problem entry point is a batch of one
  ConstrainedNonlinearLeastSquares nls(&problem);
  const NLSSolverOutputs one = nls.Solve(p, guesses[1]);
  CHECK(one.termination_state == out[1].termination_state && one.iterations.size() == out[1].iterations.size());
  CHECK(nls.variables()[0] == out[1].variables[0] && nls.variables()[1] == out[1].variables[1]);
  // SetUserExitCallback (nonlinear.hpp:150-153, nonlinear.cc:145-151): returning false after the first iteration ends the
  // solve with USER_CALLBACK and exactly one logged iteration; returning true changes nothing
  int calls = 0;
  nls.SetUserExitCallback([&calls](const NLSIteration& it) { ++calls; return it.iteration < 0; });
  const NLSSolverOutputs stopped = nls.Solve(p, guesses[1]);
  CHECK(stopped.termination_state == NLSTerminationState::USER_CALLBACK && stopped.iterations.size() == 1 && calls == 1);
  nls.SetUserExitCallback([&calls](const NLSIteration&) { ++calls; return true; });
  const NLSSolverOutputs again = nls.Solve(p, guesses[1]);
  CHECK(again.termination_state == one.termination_state && again.iterations.size() == one.iterations.size());
  CHECK(calls == 1 + static_cast<int>(one.iterations.size()));
}

static void TestInequalityConstrainedRosenbrock() {  // nonlinear_test.cc:463-500
  Problem problem;
  problem.costs.push_back(MakeResidual<2, 2>({0, 1}, &Rosenbrock));
  problem.dimension = 2;
  problem.inequality_constraints.push_back(Var(0) >= 1.2);
  problem.inequality_constraints.push_back(Var(1) <= 0.5);
  ConstrainedNonlinearLeastSquares::Params p;
  p.max_iterations = 10;
  p.max_qp_iterations = 10;
  const std::vector<Vector> guesses = {{12, -5}, {100.0, -20.0}, {1423.0, -400.0}, {-20.0, 10.0}, {-120.0, 35.0}, {-50.0, 0.5}};
  const std::vector<const Problem*> problems(guesses.size(), &problem);
  const std::vector<NLSSolverOutputs> out = ConstrainedNonlinearLeastSquares::SolveBatch(problems, p, guesses);
  PrintResults("rosenbrock_ineq", out);
  for (const NLSSolverOutputs& o : out) {
    CHECK(o.termination_state != NLSTerminationState::MAX_ITERATIONS && o.termination_state != NLSTerminationState::MAX_LAMBDA);
    CHECK(std::abs(o.variables[0] - 1.2) < 1e-6 && std::abs(o.variables[1] - 0.5) < 1e-6);
  }
  // per-step QP logs (NLSIteration::qp_outputs, structs.hpp:262-263) on request: same solve, same counts, printable
  p.record_qp_iterations = true;
  const std::vector<NLSSolverOutputs> logged = ConstrainedNonlinearLeastSquares::SolveBatch(problems, p, guesses);
  for (std::size_t i = 0; i < out.size(); ++i) {
    CHECK(logged[i].termination_state == out[i].termination_state && logged[i].iterations.size() == out[i].iterations.size());
    // asking for the log may select another kernel family for the inner solves (same algorithm, different rounding): the
    // start (1423, -400) with G ~ 8e8 moves by up to two Newton steps, see tests/test_cpp_host_api.py
    const long dq = static_cast<long>(logged[i].NumQPIterations()) - static_cast<long>(out[i].NumQPIterations());
    CHECK(dq >= -2 && dq <= 2 && std::abs(logged[i].variables[0] - out[i].variables[0]) < 1e-7);
    std::size_t steps = 0;
    for (const NLSIteration& it : logged[i].iterations) {
      const QPInteriorPointSolverOutputs* ip = std::get_if<QPInteriorPointSolverOutputs>(&it.qp_outputs);
      CHECK(ip != nullptr && static_cast<int>(ip->iterations.size()) == it.qp_iterations);
      steps += ip->iterations.size();
    }
    CHECK(steps == logged[i].NumQPIterations());
    const std::string text = logged[i].ToString(false, true);
    CHECK(text.find("  QP-IP: ") != std::string::npos && text.find("Iteration summary: ||kkt|| max: ") != std::string::npos);
    CHECK(logged[i].ToJson().find("\"kkt_initial\"") != std::string::npos);
  }
}

static void TestHimmelblauQuadrantConstrained() {  // nonlinear_test.cc:668-718
  Problem problem;
  problem.costs.push_back(MakeResidual<1, 2>({0, 1}, [](const std::array<double, 2>& xy, std::array<double, 2>* J) {
    if (J) { (*J)[0] = 2 * xy[0]; (*J)[1] = 1.0; }
    return std::array<double, 1>{std::pow(xy[0], 2) + xy[1] - 11};
  }));
  problem.costs.push_back(MakeResidual<1, 2>({0, 1}, [](const std::array<double, 2>& xy, std::array<double, 2>* J) {
    if (J) { (*J)[0] = 1.0; (*J)[1] = 2 * xy[1]; }
    return std::array<double, 1>{xy[0] + std::pow(xy[1], 2) - 7};
  }));
  problem.dimension = 2;
  problem.inequality_constraints = {Var(0) >= 0.1, Var(0) <= 5.0, Var(1) >= 0.1, Var(1) <= 5.0};
  ConstrainedNonlinearLeastSquares::Params p;
  p.max_iterations = 20;
  p.max_qp_iterations = 10;
  p.relative_exit_tol = 1e-12;
  p.absolute_first_derivative_tol = 1e-8;
  p.termination_kkt_tolerance = 1e-6;
  std::vector<Vector> guesses;
  for (double x = 0.2; x <= 4.8; x += 0.2)
    for (double y = 0.2; y <= 4.8; y += 0.2) guesses.push_back(Vector{x, y});
  const std::vector<const Problem*> problems(guesses.size(), &problem);
  const std::vector<NLSSolverOutputs> out = ConstrainedNonlinearLeastSquares::SolveBatch(problems, p, guesses);
  PrintResults("himmelblau_quadrant", out);
  for (const NLSSolverOutputs& o : out) {
    CHECK(SatisfiedTol(o.termination_state));
    CHECK(std::abs(o.variables[0] - 3.0) < 5e-5 && std::abs(o.variables[1] - 2.0) < 5e-5);  // the reference's own bound
  }
}

static void TestSphereWithNonlinearEqualityConstraints() {  // nonlinear_test.cc:745-830
  constexpr int N = 6;
  Problem problem;
  problem.costs.push_back(MakeResidual<N, N>({0, 1, 2, 3, 4, 5}, &SphereFunction<N>));
  problem.dimension = N;
  problem.equality_constraints.push_back(MakeResidual<1, 2>(
      {0, 1}, [](const std::array<double, 2>& x, std::array<double, 2>* J) { return ProductExpression(x, 4.0, J); }));
  problem.equality_constraints.push_back(MakeResidual<1, 2>(
      {2, 3}, [](const std::array<double, 2>& x, std::array<double, 2>* J) { return ProductExpression(x, 9.0, J); }));
  ConstrainedNonlinearLeastSquares::Params p;
  p.max_iterations = 100;
  p.max_qp_iterations = 1;
  p.relative_exit_tol = 1e-12;
  p.absolute_first_derivative_tol = 1e-9;
  p.termination_kkt_tolerance = 1e-6;
  p.lambda_initial = 0.001;
  std::vector<Vector> guesses;
  std::mt19937 engine{7};
  std::uniform_real_distribution<double> dist{-30.0, 30.0};
  for (int i = 0; i < 100; ++i) {
    Vector g(N);
    for (int j = 0; j < N; ++j) g[j] = dist(engine);
    guesses.push_back(g);
  }
  const std::vector<const Problem*> problems(guesses.size(), &problem);
  const std::vector<NLSSolverOutputs> out = ConstrainedNonlinearLeastSquares::SolveBatch(problems, p, guesses);
  PrintResults("sphere", out);
  int good = 0;
  for (const NLSSolverOutputs& o : out) {
    CHECK(SatisfiedTol(o.termination_state));
    const Vector& v = o.variables;
    // one of the four solutions (+-2, +-2, +-3, +-3, 0, 0)
    const bool ok = std::abs(std::abs(v[0]) - 2.0) < 1e-4 && std::abs(v[0] - v[1]) < 1e-4 && std::abs(std::abs(v[2]) - 3.0) < 1e-4 &&
                    std::abs(v[2] - v[3]) < 1e-4 && std::abs(v[4]) < 1e-4 && std::abs(v[5]) < 1e-4;
    good += ok;
  }
  CHECK(good == static_cast<int>(out.size()));
}

// A singular problem inside a batch ends as QP_FAILED for itself; the others go on. The batch-of-one
// Solve() throws FailedFactorization like the reference (qp.cc:303-307). CheckParams follows nonlinear.cc:49-73.
static void TestPerProblemFailureAndParams() {
  Problem good;
  good.dimension = 2;
  good.costs.push_back(MakeResidual<2, 2>({0, 1}, &Rosenbrock));
  good.inequality_constraints.push_back(Var(0) >= -50.0);
  Problem singular;  // J = 0 and lambda = 0: G = 0 -> zero pivots in the KKT factorization
  singular.dimension = 2;
  singular.costs.push_back(MakeResidual<2, 2>({0, 1}, [](const std::array<double, 2>&, std::array<double, 4>* J) {
    if (J) J->fill(0.0);
    return std::array<double, 2>{1.0, 1.0};
  }));
  singular.inequality_constraints.push_back(Var(0) >= -50.0);
  ConstrainedNonlinearLeastSquares::Params p;
  p.max_iterations = 20;
  p.max_qp_iterations = 10;
  const std::vector<NLSSolverOutputs> out =
      ConstrainedNonlinearLeastSquares::SolveBatch({&good, &singular, &good}, p, {Vector{0.8, -0.3}, Vector{0.0, 0.0}, Vector{4.0, 0.0}});
  CHECK(out[1].termination_state == NLSTerminationState::QP_FAILED && out[1].qp_failure_status == MO_QP_FAILED_FACTORIZATION);
  CHECK(SatisfiedTol(out[0].termination_state) && SatisfiedTol(out[2].termination_state));
  CHECK(std::abs(out[0].variables[0] - 1.0) < 1e-5 && std::abs(out[2].variables[1] - 1.0) < 1e-5);
  ConstrainedNonlinearLeastSquares one(&singular);
  bool threw = false;
  try { (void)one.Solve(p, Vector{0.0, 0.0}); } catch (const FailedFactorization&) { threw = true; }
  CHECK(threw);
  // values the reference accepts (>= 0) and values it asserts against
  ConstrainedNonlinearLeastSquares nls(&good);
  ConstrainedNonlinearLeastSquares::Params q = p;
  q.equality_penalty_initial = 0.0; q.lambda_failure_init = 0.0;
  (void)nls.Solve(q, Vector{0.8, -0.3});
  for (int bad = 0; bad < 4; ++bad) {
    q = p;
    if (bad == 0) { q.min_lambda = 2.0; q.max_lambda = 1.0; }
    if (bad == 1) q.lambda_decrease_on_success = 1.0;
    if (bad == 2) q.equality_penalty_scale_factor = 0.5;
    if (bad == 3) q.equality_penalty_rho = 1.0;
    threw = false;
    try { (void)nls.Solve(q, Vector{0.8, -0.3}); } catch (const std::invalid_argument&) { threw = true; }
    CHECK(threw);
  }
}

int main(int argc, char** argv) {
  const std::string mode = argc > 1 ? argv[1] : "cpu";
  TestLinearization();
  if (mode == "cpu") { std::printf("cpu ok\n"); return 0; }
  TestRosenbrock(false);
  TestRosenbrock(true);
  TestInequalityConstrainedRosenbrock();
  TestHimmelblauQuadrantConstrained();
  TestSphereWithNonlinearEqualityConstraints();
  TestPerProblemFailureAndParams();
  std::printf("gpu ok\n");
  return 0;
}
