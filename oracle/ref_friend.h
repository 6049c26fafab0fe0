// TEST INFRASTRUCTURE — not part of the product.
// qp.hpp:293-294 and nonlinear.hpp:245 befriend a class named `ConstrainedNLSTest` (the reference's own test
// fixture). The reference's test binaries define it; the harness defines it here, once, to reach the private
// members it needs for step-level parity: QPInteriorPointSolver::Iterate / delta_ and
// ConstrainedNonlinearLeastSquares::LinearizeAndFillQP.
#pragma once
#include <Eigen/Dense>

#include "mini_opt/nonlinear.hpp"
#include "mini_opt/qp.hpp"

namespace mini_opt {

class ConstrainedNLSTest {
 public:
  static IPIterationOutputs Iterate(QPInteriorPointSolver& s, double mu, BarrierStrategy strat) {
    return s.Iterate(mu, strat);
  }
  static const Eigen::VectorXd& Delta(const QPInteriorPointSolver& s) { return s.delta_; }
  static void EvaluateKKT(QPInteriorPointSolver& s) { s.EvaluateKKTConditions(); }
  static KKTError Errors(const QPInteriorPointSolver& s, double mu) { return s.ComputeErrors(mu); }
  static Eigen::VectorXd& Vars(QPInteriorPointSolver& s) { return s.variables_; }
  static mini_opt::Errors Linearize(const Eigen::VectorXd& x, double lambda, const Problem& problem, QP* qp) {
    return ConstrainedNonlinearLeastSquares::LinearizeAndFillQP(x, lambda, problem, qp);
  }
};

}  // namespace mini_opt
