// C++ host API smoke test (built and run by tests/test_cpp_host_api.py). Mirrors the reference's own unit tests:
//   argv[1] == "cpu": no device needed -- packing, Setup assertions (qp.cc:20-44), CheckParams (qp.cc:76-82)
//   argv[1] == "gpu": TestInequalityConstraints-style 1-D problem (x -> 4, test/qp_test.cc:252-288) and a batch
#include <cmath>
#include <cstdio>
#include <cstring>

#include "mini_opt_b200/qp.hpp"

using namespace mini_opt_b200;
#define CHECK(cond) do { if (!(cond)) { std::printf("CHECK failed line %d: %s\n", __LINE__, #cond); return 1; } } while (0)

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  QP qp(1);  // (x - 5)^2 = x^2 - 10 x + 25 -> G = 2, c = -10; x <= 4
  qp.G(0, 0) = 2.0;
  qp.c[0] = -10.0;
  qp.constraints.push_back(Var(0) <= 4.0);
  CHECK(qp.constraints[0].a == -1.0 && qp.constraints[0].b == 4.0 && qp.constraints[0].IsFeasible(3.0));
  CHECK((Var(0) >= 1.5).ClampX(0.0) == 1.5);
  QPInteriorPointSolver solver(&qp);
  QPInteriorPointSolver::Params params{};
  params.max_iterations = 0;  // CheckParams, qp.cc:76-82
  bool threw = false;
  try { (void)solver.Solve(params); } catch (const std::invalid_argument&) { threw = true; }
  CHECK(threw);
  QP bad(2);
  bad.constraints.push_back(Var(5) <= 1.0);  // index out of bounds, qp.cc:36-40
  threw = false;
  try { QPInteriorPointSolver s2(&bad); } catch (const std::invalid_argument&) { threw = true; }
  CHECK(threw);
  if (!gpu) {  // without a device the solve must fail loudly: there is no CPU fallback
    params.max_iterations = 10;
    threw = false;
    try { (void)solver.Solve(params); } catch (const std::runtime_error&) { threw = true; }
    CHECK(threw);
    std::printf("cpu ok\n");
    return 0;
  }
  params = QPInteriorPointSolver::Params{};
  params.max_iterations = 30;
  params.sigma = 0.1;
  const QPInteriorPointSolverOutputs out = solver.Solve(params);
  CHECK(out.termination_state == QPInteriorPointTerminationState::SATISFIED_KKT_TOL);
  CHECK(std::fabs(out.x(0) - 4.0) < 1e-6 && std::fabs(solver.variables()[0] - 4.0) < 1e-6);
  CHECK(std::fabs(out.z(0) - 2.0) < 1e-5);  // multiplier of the active bound: G x + c + z = 0 -> z = 2
  // the per-step log of the reference's outputs (structs.hpp:120-122), from iter_log
  CHECK(static_cast<int>(out.iterations.size()) == out.num_iterations && out.num_iterations > 0);
  CHECK(out.iterations.back().kkt_final.Max() == out.kkt_final.Max() && out.kkt_final.Max() < 1e-9);
  CHECK(out.iterations.front().ip_outputs.mu == 1.0 && std::isnan(out.iterations.front().ip_outputs.mu_affine));
  for (std::size_t i = 1; i < out.iterations.size(); ++i)  // kkt_initial of a step is evaluated with the new mu, its r_dual is the previous step's
    CHECK(std::fabs(out.iterations[i].kkt_initial.r_dual - out.iterations[i - 1].kkt_final.r_dual) <= 1e-12 * (1.0 + out.iterations[i].kkt_initial.r_dual));
  CHECK(out.iterations[0].ToString().rfind("Iteration summary: ||kkt|| max: ", 0) == 0);
  CHECK(out.ToJson().find("\"termination_state\":\"SATISFIED_KKT_TOL\"") != std::string::npos);
  std::vector<QP> batch;
  for (int i = 0; i < 100; ++i) {
    QP q(2);
    q.G(0, 0) = 2.0; q.G(1, 1) = 1.0 + i * 0.01; q.c[0] = -10.0; q.c[1] = 3.0;
    q.constraints = {Var(0) >= -1.0, Var(0) <= 4.0, Var(1) >= -1.0, Var(1) <= 1.0};
    batch.push_back(q);
  }
  const auto res = QPInteriorPointSolver::SolveBatch(batch, params);
  for (const auto& o : res) CHECK(o.batch_status == 0 && std::fabs(o.x(0) - 4.0) < 1e-6 && std::fabs(o.x(1) + 1.0) < 1e-6);
  // USER_PROVIDED through SetVariables (qp.cc:221): start at the solved point -> converges again, in few iterations
  {
    QPInteriorPointSolver::Params up = params;
    up.initial_guess_method = InitialGuessMethod::USER_PROVIDED;
    QPInteriorPointSolver cold(&qp);
    threw = false;
    try { (void)cold.Solve(up); } catch (const std::invalid_argument&) { threw = true; }  // no SetVariables yet
    CHECK(threw);
    Vector start = {3.0, 1.0, 1.0};  // [x | s | z]: interior point
    cold.SetVariables(start);
    const QPInteriorPointSolverOutputs warm = cold.Solve(up);
    CHECK(warm.termination_state == QPInteriorPointTerminationState::SATISFIED_KKT_TOL && std::fabs(warm.x(0) - 4.0) < 1e-6);
    CHECK(cold.x_block().size() == 1 && cold.s_block().size() == 1 && cold.y_block().empty() && cold.z_block().size() == 1);
    CHECK(std::fabs(cold.x_block()[0] - 4.0) < 1e-6 && std::fabs(cold.z_block()[0] - 2.0) < 1e-5);
    threw = false;
    try { cold.SetVariables(Vector{1.0}); } catch (const std::invalid_argument&) { threw = true; }
    CHECK(threw);
  }
  // QPNullSpaceSolver: TestNullSpaceSolver1 (test/qp_test.cc:576-626): |A x - b|^2 with x1 pinned to 0.3
  {
    QP ns(2);
    const double A[2][2] = {{-2.0, 1.4}, {2.2, -3.5}}, b[2] = {-0.8, 1.3};
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 2; ++j) {
        ns.G(i, j) = A[0][i] * A[0][j] + A[1][i] * A[1][j];  // UpdateHessian: J'J
      }
    for (int i = 0; i < 2; ++i) ns.c[i] = -(A[0][i] * b[0] + A[1][i] * b[1]);  // J'r at x = 0
    ns.A_eq.resize(1, 2);
    ns.A_eq(0, 1) = 1.0;
    ns.b_eq = Vector{-0.3};
    QPNullSpaceSolver nss;
    CHECK(nss.Solve(ns) == QPNullSpaceTerminationState::SUCCESS);
    CHECK(std::fabs(nss.variables()[1] - 0.3) < 1e-12 && std::fabs(nss.variables()[0] - 0.860859728506787) < 1e-12);
    QP flat = ns;
    flat.G = Matrix(2, 2);  // reduced Hessian 0 -> NOT_POSITIVE_DEFINITE
    CHECK(nss.Solve(flat) == QPNullSpaceTerminationState::NOT_POSITIVE_DEFINITE);
  }
  std::printf("gpu ok\n");
  return 0;
}
