// TEST INFRASTRUCTURE. Prints what the UNMODIFIED reference's ToString() (source/structs.cc:122-268) and JSON serializer
// (source/serialization.cc:65-134, nlohmann::json) produce for a fixed set of hand-written struct values. The same values are
// built by tests/cpp/test_logging.cc with the structs of include/mini_opt_b200/{qp,nonlinear}.hpp; tests/test_cpp_host_api.py
// compares the text blocks byte for byte and the JSON documents as parsed objects.
// Build + run: make -C oracle log_golden   (writes tests/golden/logging_reference.txt; needs /root/reference)
#include <cstdio>
#include <limits>

#include <nlohmann/json.hpp>

#include "mini_opt/serialization.hpp"
#include "mini_opt/structs.hpp"

using namespace mini_opt;

static QPInteriorPointIteration QpIter(int i, bool probe) {
  IPIterationOutputs ip{};
  ip.mu = 1.0 / (3.0 + i);
  ip.alpha = AlphaValues{0.995 / (1 + i), 1.0};
  if (probe) {
    ip.alpha_probe = AlphaValues{0.25, 1e-5};
    ip.mu_affine = 123456789.125;
  }
  return QPInteriorPointIteration(KKTError{1e-9 * (i + 1), 0.1, 0.0, 2.5e16}, KKTError{1e-3 / 3.0, 1e-5, 100000.0, 1e15}, ip);
}

int main() {
  const QPInteriorPointIteration a = QpIter(0, false), b = QpIter(1, true);
  std::printf("### QP_ITER_PLAIN\n%s### QP_ITER_PROBE\n%s", a.ToString().c_str(), b.ToString().c_str());

  QPInteriorPointSolverOutputs ip_out(QPInteriorPointTerminationState::SATISFIED_KKT_TOL, {a, b}, QPLagrangeMultipliers{-0.5, 2.0});
  QPInteriorPointSolverOutputs ip_out2(QPInteriorPointTerminationState::MAX_ITERATIONS, {b}, std::nullopt);
  std::vector<LineSearchStep> steps = {LineSearchStep{1.0, Errors{0.75, 1e-3}}, LineSearchStep{0.5, Errors{0.0625, 1e-12}}};
  NLSIteration it0(0, OptimizerState::NOMINAL, 0.0, Errors{12.5, 0.3}, ip_out, std::nullopt, DirectionalDerivatives{-3.25, -0.3},
                   1.01, StepSizeSelectionResult::SUCCESS, steps);
  NLSIteration it1(1, OptimizerState::ATTEMPTING_RESTORE_LM, 1e-3, Errors{0.0625, 1e-12}, QPNullSpaceTerminationState::SUCCESS,
                   std::nullopt, DirectionalDerivatives{-1e-7, 0.0}, 2.0 / 3.0, StepSizeSelectionResult::MAX_ITERATIONS, {});
  NLSIteration it2(2, OptimizerState::NOMINAL, 10.0, Errors{1e-20, 0.0}, ip_out2, std::nullopt, DirectionalDerivatives{0.0, 0.0},
                   1e6, StepSizeSelectionResult::FIRST_ORDER_SATISFIED, {LineSearchStep{0.125, Errors{1e-20, 0.0}}});
  NLSSolverOutputs out(NLSTerminationState::SATISFIED_FIRST_ORDER_TOL, {it0, it1, it2});
  std::printf("### NLS_PLAIN\n%s", out.ToString(false, false).c_str());
  std::printf("### NLS_COLOR_QP\n%s", out.ToString(true, true).c_str());
  std::printf("### NLS_COUNTS\n%zu %zu %zu\n", out.NumQPIterations(), out.NumLineSearchSteps(), out.NumFailedLineSearches());
  const nlohmann::json j = out;
  std::printf("### NLS_JSON\n%s\n### END\n", j.dump().c_str());  // QPInteriorPointSolverOutputs appears as it0.qp_outputs
  return 0;
}
