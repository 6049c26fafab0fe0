// TEST INFRASTRUCTURE — not part of the product (see ref_harness.cc).
//
// Exposes the reference's OWN synthetic QP generator, QPSolverTest::GenerateRandomQP
// (test/qp_test.cc:483-524, which uses test_utils::GenerateRandomPDMatrix, test/test_utils.cc:19-34),
// by compiling the reference's test translation unit as-is and deriving from its fixture. The generator
// draws from std::default_random_engine / std::uniform_real_distribution, whose streams are
// libstdc++-specific, so the only way to get the reference's exact problems is to run its code; the
// vectors are then frozen under tests/golden/ by tools/make_golden.py.
#include <cstring>

// The reference's test TU (path supplied by oracle/Makefile via -I/root/reference/test).
#include "qp_test.cc"

namespace {
struct Generator final : public mini_opt::QPSolverTest {
  void TestBody() override {}
};
}  // namespace

extern "C" {

// Writes problem `seed` into caller buffers: G [dim*dim] col-major, c [dim], constraint triplets
// (var, a, b) with capacity dim, solution [dim]. Returns the number of constraints.
int ref_generate_random_qp(int seed, int dim, double p_inequality, double* G, double* c, int32_t* var,
                           double* a, double* b, double* solution) {
  Generator gen;
  Eigen::VectorXd sol;
  std::vector<uint8_t> mask;
  const mini_opt::QP qp = gen.GenerateRandomQP(seed, dim, p_inequality, &sol, &mask);
  std::memcpy(G, qp.G.data(), sizeof(double) * dim * dim);
  std::memcpy(c, qp.c.data(), sizeof(double) * dim);
  std::memcpy(solution, sol.data(), sizeof(double) * dim);
  int count = 0;
  for (const auto& con : qp.constraints) {
    var[count] = con.variable;
    a[count] = con.a;
    b[count] = con.b;
    ++count;
  }
  return count;
}

}  // extern "C"
