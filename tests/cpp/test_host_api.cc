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
  // General inequalities through the exact lifting (SolveBatchWithGeneralInequalities; the reference has none, README.md:36-37):
  // min (x-2)^2 + (y-2)^2  s.t.  x + y <= 2, x >= 0, y >= 0  ->  (1, 1), multiplier of the general constraint 2
  {
    QP tri(2);
    tri.G(0, 0) = 2.0; tri.G(1, 1) = 2.0; tri.c[0] = -4.0; tri.c[1] = -4.0;
    tri.constraints = {Var(0) >= 0.0, Var(1) >= 0.0};
    const std::vector<GeneralInequality> gen = {GeneralInequality{Vector{-1.0, -1.0}, 2.0}};
    QPInteriorPointSolver::Params gp;
    gp.max_iterations = 40;
    gp.sigma = 0.1;
    const auto r1 = SolveBatchWithGeneralInequalities({tri}, {gen}, gp);
    CHECK(r1[0].termination_state == QPInteriorPointTerminationState::SATISFIED_KKT_TOL && r1[0].n == 2 && r1[0].m == 3 && r1[0].k == 0);
    CHECK(std::fabs(r1[0].x(0) - 1.0) < 1e-6 && std::fabs(r1[0].x(1) - 1.0) < 1e-6);
    CHECK(std::fabs(r1[0].z(2) - 2.0) < 1e-5 && std::fabs(r1[0].s(2)) < 1e-6 && r1[0].z(0) < 1e-6);
    // a batch of random 6-variable problems, 1 equality row, box on every variable, 4 general constraints: the KKT
    // conditions of the ORIGINAL problem hold at the returned point
    std::vector<QP> qps;
    std::vector<std::vector<GeneralInequality>> gens;
    unsigned long long st = 12345;
    auto rnd = [&st]() { st = st * 6364136223846793005ULL + 1442695040888963407ULL; return (double)((st >> 11) & 0xFFFFFFFFFFFFFULL) / 4503599627370496.0 * 2.0 - 1.0; };
    for (int p = 0; p < 64; ++p) {
      const int n = 6;
      QP q(n);
      double Mx[6][6];
      for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) Mx[i][j] = rnd();
      for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) {
          double acc = i == j ? 1.0 : 0.0;
          for (int l = 0; l < n; ++l) acc += Mx[i][l] * Mx[j][l];
          q.G(i, j) = acc; q.G(j, i) = acc;
        }
      for (int i = 0; i < n; ++i) q.c[i] = 3.0 * rnd();
      q.A_eq.resize(1, n);
      for (int j = 0; j < n; ++j) q.A_eq(0, j) = rnd();
      q.b_eq = Vector{0.1 * rnd()};
      for (int i = 0; i < n; ++i) { q.constraints.push_back(Var(i) >= -1.0); q.constraints.push_back(Var(i) <= 1.0); }
      std::vector<GeneralInequality> g;
      for (int j = 0; j < 4; ++j) {
        GeneralInequality gi{Vector(n, 0.0), 0.5 + 0.25 * rnd()};
        for (int i = 0; i < n; ++i) gi.a[i] = rnd();
        g.push_back(gi);
      }
      qps.push_back(q);
      gens.push_back(g);
    }
    const auto rb = SolveBatchWithGeneralInequalities(qps, gens, gp);
    int converged = 0;
    for (std::size_t p = 0; p < qps.size(); ++p) {
      const QPInteriorPointSolverOutputs& o = rb[p];
      if (o.termination_state != QPInteriorPointTerminationState::SATISFIED_KKT_TOL) continue;
      ++converged;
      CHECK(o.n == 6 && o.k == 1 && o.m == 16);
      Vector x(6);
      for (int i = 0; i < 6; ++i) x[i] = o.x(i);
      for (int i = 0; i < 6; ++i) {  // stationarity: G x + c - A' y - sum a_i z_i (diagonal: a = +-1 on variable i/2; general: dense)
        double rsd = qps[p].c[i] - qps[p].A_eq(0, i) * o.y(0);
        for (int j = 0; j < 6; ++j) rsd += qps[p].G(i, j) * x[j];
        rsd -= o.z(2 * i) - o.z(2 * i + 1);
        for (int j = 0; j < 4; ++j) rsd -= gens[p][j].a[i] * o.z(12 + j);
        CHECK(std::fabs(rsd) < 1e-6);
      }
      double eq = qps[p].b_eq[0];
      for (int j = 0; j < 6; ++j) eq += qps[p].A_eq(0, j) * x[j];
      CHECK(std::fabs(eq) < 1e-7);
      for (int j = 0; j < 4; ++j) {
        const double v = gens[p][j].Evaluate(x);
        CHECK(v > -1e-7 && o.z(12 + j) >= 0.0 && std::fabs(v * o.z(12 + j)) < 1e-5 && std::fabs(v - o.s(12 + j)) < 1e-7);
      }
      for (int i = 0; i < 6; ++i) CHECK(x[i] > -1.0 - 1e-7 && x[i] < 1.0 + 1e-7);
    }
    CHECK(converged >= 62);
  }
  std::printf("gpu ok\n");
  return 0;
}
