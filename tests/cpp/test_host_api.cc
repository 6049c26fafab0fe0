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
  std::vector<QP> batch;
  for (int i = 0; i < 100; ++i) {
    QP q(2);
    q.G(0, 0) = 2.0; q.G(1, 1) = 1.0 + i * 0.01; q.c[0] = -10.0; q.c[1] = 3.0;
    q.constraints = {Var(0) >= -1.0, Var(0) <= 4.0, Var(1) >= -1.0, Var(1) <= 1.0};
    batch.push_back(q);
  }
  const auto res = QPInteriorPointSolver::SolveBatch(batch, params);
  for (const auto& o : res) CHECK(o.batch_status == 0 && std::fabs(o.x(0) - 4.0) < 1e-6 && std::fabs(o.x(1) + 1.0) < 1e-6);
  std::printf("gpu ok\n");
  return 0;
}
