// CPU-only: with Eigen on the include path a residual functor written for the REFERENCE (Eigen fixed-size signature,
// residual.hpp:121-143; here the reference test's Rosenbrock, nonlinear_test.cc:375-387, restated) goes through
// MakeResidualEigen and linearizes to the same QP as the std::array form. Built by tests/test_cpp_host_api.py only when an
// Eigen tree is available (this container: /root/reference/dependencies/eigen; the GPU box has none and skips).
#define MINI_OPT_B200_WITH_EIGEN
#include <cmath>
#include <cstdio>

#include "mini_opt_b200/nonlinear.hpp"

using namespace mini_opt_b200;

static Eigen::Vector2d RosenbrockEigen(const Eigen::Vector2d& p, Eigen::Matrix2d* J) {
  const double sb = std::sqrt(100.0);
  if (J) { J->setZero(); (*J)(0, 0) = -1.0; (*J)(1, 0) = -2 * p.x() * sb; (*J)(1, 1) = sb; }
  return {1.0 - p.x(), sb * (p.y() - p.x() * p.x())};
}
static std::array<double, 2> RosenbrockArray(const std::array<double, 2>& p, std::array<double, 4>* J) {
  const double sb = std::sqrt(100.0);
  if (J) { (*J)[0] = -1.0; (*J)[1] = -2 * p[0] * sb; (*J)[2] = 0.0; (*J)[3] = sb; }
  return {1.0 - p[0], sb * (p[1] - p[0] * p[0])};
}

int main() {
  Problem a, b;
  a.dimension = b.dimension = 3;
  a.costs.push_back(MakeResidualEigen<2, 2>({2, 0}, &RosenbrockEigen));
  b.costs.push_back(MakeResidual<2, 2>({2, 0}, &RosenbrockArray));
  const Vector x = FromEigen(Eigen::Vector3d(0.5, -1.5, 2.0));
  QP qa, qb;
  const Errors ea = ConstrainedNonlinearLeastSquares::LinearizeAndFillQP(x, 0.25, a, &qa);
  const Errors eb = ConstrainedNonlinearLeastSquares::LinearizeAndFillQP(x, 0.25, b, &qb);
  if (ea.f != eb.f || ToEigen(x).size() != 3) { std::printf("errors differ\n"); return 1; }
  for (int i = 0; i < 3; ++i) {
    if (qa.c[i] != qb.c[i]) { std::printf("c differs\n"); return 1; }
    for (int j = 0; j < 3; ++j)
      if (qa.G(i, j) != qb.G(i, j)) { std::printf("G differs\n"); return 1; }
  }
  std::printf("eigen ok\n");
  return 0;
}
