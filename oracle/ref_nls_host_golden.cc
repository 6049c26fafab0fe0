// TEST INFRASTRUCTURE: runs the UNMODIFIED reference's ConstrainedNonlinearLeastSquares on the reference's own
// nonlinear test problems (gareth-cross/mini_opt test/nonlinear_test.cc:375-500: Rosenbrock, Rosenbrock with LM only,
// inequality constrained Rosenbrock; the three residual functions are restated here because they are private members
// of the reference's test fixture) and prints one line per start point:
//   <test> <index> <termination_state> <outer iterations> <total QP iterations> <x...>
// tools/make_golden_nls_host.sh links this against oracle/_ref/libminiopt_ref.so and writes
// tests/golden/nls_host_reference.txt, which tests/test_cpp_host_api.py compares with the output of
// include/mini_opt_b200/nonlinear.hpp's SolveBatch on the same problems.
#include <cmath>
#include <cstdio>
#include <random>
#include <vector>

#include "mini_opt/nonlinear.hpp"
#include "mini_opt/residual.hpp"

using namespace mini_opt;
using Eigen::Matrix2d;
using Eigen::Vector2d;

static Vector2d Rosenbrock(const Vector2d& params, Matrix2d* const J_out) {
  const double sb = std::sqrt(100.0);
  if (J_out) {
    J_out->setZero();
    (*J_out)(0, 0) = -1.0;
    (*J_out)(1, 0) = -2 * params.x() * sb;
    (*J_out)(1, 1) = sb;
  }
  return {1.0 - params.x(), sb * (params.y() - params.x() * params.x())};
}

static void Run(const char* name, const Problem& problem, const ConstrainedNonlinearLeastSquares::Params& p,
                const std::vector<Vector2d>& guesses) {
  ConstrainedNonlinearLeastSquares nls(&problem);
  for (std::size_t i = 0; i < guesses.size(); ++i) {
    const NLSSolverOutputs out = nls.Solve(p, guesses[i]);
    std::printf("%s %zu %d %zu %d %.17g %.17g\n", name, i, static_cast<int>(out.termination_state), out.iterations.size(),
                out.NumQPIterations(), nls.variables()[0], nls.variables()[1]);
  }
}

int main() {
  const std::vector<Vector2d> guesses = {{-5, -3}, {10, 8}, {-20, 3}, {0, -5}, {4, 0}, {100, 50}, {-35, 40}, {1000, -50}, {0.8, -0.3}};
  {
    Problem problem{};
    problem.costs.push_back(MakeResidual<2, 2>({0, 1}, &Rosenbrock));
    problem.dimension = 2;
    ConstrainedNonlinearLeastSquares::Params p{};
    p.max_iterations = 5;
    p.max_qp_iterations = 1;
    Run("rosenbrock", problem, p, guesses);
    p.max_iterations = 10;
    p.absolute_first_derivative_tol = 1e-12;
    p.max_line_search_iterations = 0;
    Run("rosenbrock_lm", problem, p, guesses);
  }
  {
    Problem problem{};
    problem.costs.emplace_back(MakeResidual<2, 2>({0, 1}, &Rosenbrock));
    problem.dimension = 2;
    problem.inequality_constraints.push_back(Var(0) >= 1.2);
    problem.inequality_constraints.push_back(Var(1) <= 0.5);
    ConstrainedNonlinearLeastSquares::Params p{};
    p.max_iterations = 10;
    p.max_qp_iterations = 10;
    Run("rosenbrock_ineq", problem, p,
        {{12, -5}, {100.0, -20.0}, {1423.0, -400.0}, {-20.0, 10.0}, {-120.0, 35.0}, {-50.0, 0.5}});
  }
  {
    // TestSphereWithNonlinearEqualityConstraints (nonlinear_test.cc:745-830): n = 6, two nonlinear equalities, no
    // inequalities -> QPNullSpaceSolver + the (18.36) penalty. 100 seeded start points (std::mt19937{7}, as in
    // tests/cpp/test_nls_host_api.cc); x0, x1 of the result are printed.
    constexpr int N = 6;
    Problem problem{};
    problem.costs.push_back(MakeResidual<N, N>({0, 1, 2, 3, 4, 5}, [](const Eigen::Matrix<double, N, 1>& x, Eigen::Matrix<double, N, N>* J) {
      if (J) J->setIdentity();
      return Eigen::Matrix<double, N, 1>(x);
    }));
    problem.dimension = N;
    auto product = [](double v) {
      return [v](const Vector2d& x, Eigen::Matrix<double, 1, 2>* J) -> Eigen::Matrix<double, 1, 1> {
        if (J) { (*J)(0, 0) = x[1]; (*J)(0, 1) = x[0]; }
        return Eigen::Matrix<double, 1, 1>{x[0] * x[1] - v};
      };
    };
    problem.equality_constraints.push_back(MakeResidual<1, 2>({0, 1}, product(4.0)));
    problem.equality_constraints.push_back(MakeResidual<1, 2>({2, 3}, product(9.0)));
    ConstrainedNonlinearLeastSquares::Params p{};
    p.max_iterations = 100;
    p.max_qp_iterations = 1;
    p.relative_exit_tol = 1e-12;
    p.absolute_first_derivative_tol = 1e-9;
    p.termination_kkt_tolerance = 1e-6;
    p.lambda_initial = 0.001;
    std::mt19937 engine{7};
    std::uniform_real_distribution<double> dist{-30.0, 30.0};
    ConstrainedNonlinearLeastSquares nls(&problem);
    for (int i = 0; i < 100; ++i) {
      Eigen::VectorXd g(N);
      for (int j = 0; j < N; ++j) g[j] = dist(engine);
      const NLSSolverOutputs out = nls.Solve(p, g);
      std::printf("sphere %d %d %zu %d %.17g %.17g\n", i, static_cast<int>(out.termination_state), out.iterations.size(),
                  static_cast<int>(out.NumQPIterations()), nls.variables()[0], nls.variables()[1]);
    }
  }
  {
    // TestHimmelblauQuadrantConstrained (nonlinear_test.cc:668-718): two one-row costs, box 0.1 <= x, y <= 5, start points
    // on the reference's own grid (same accumulating loop); every start must reach (3, 2)
    Problem problem{};
    problem.costs.push_back(MakeResidual<1, 2>({0, 1}, [](const Vector2d& xy, Eigen::Matrix<double, 1, 2>* J) -> Eigen::Matrix<double, 1, 1> {
      if (J) { (*J)(0, 0) = 2 * xy[0]; (*J)(0, 1) = 1.0; }
      return Eigen::Matrix<double, 1, 1>{std::pow(xy[0], 2) + xy[1] - 11};
    }));
    problem.costs.push_back(MakeResidual<1, 2>({0, 1}, [](const Vector2d& xy, Eigen::Matrix<double, 1, 2>* J) -> Eigen::Matrix<double, 1, 1> {
      if (J) { (*J)(0, 0) = 1.0; (*J)(0, 1) = 2 * xy[1]; }
      return Eigen::Matrix<double, 1, 1>{xy[0] + std::pow(xy[1], 2) - 7};
    }));
    problem.dimension = 2;
    problem.inequality_constraints.push_back(Var(0) >= 0.1);
    problem.inequality_constraints.push_back(Var(0) <= 5.0);
    problem.inequality_constraints.push_back(Var(1) >= 0.1);
    problem.inequality_constraints.push_back(Var(1) <= 5.0);
    ConstrainedNonlinearLeastSquares::Params p{};
    p.max_iterations = 20;
    p.max_qp_iterations = 10;
    p.relative_exit_tol = 1e-12;
    p.absolute_first_derivative_tol = 1e-8;
    p.termination_kkt_tolerance = 1e-6;
    std::vector<Vector2d> grid;
    for (double x = 0.2; x <= 4.8; x += 0.2)
      for (double y = 0.2; y <= 4.8; y += 0.2) grid.emplace_back(x, y);
    Run("himmelblau_quadrant", problem, p, grid);
  }
  return 0;
}
