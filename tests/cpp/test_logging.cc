// ToString() / ToJson() of the batched result structs against the UNMODIFIED reference's output for the same values
// (tests/golden/logging_reference.txt, written by oracle/ref_log_golden.cc: structs.cc:122-268, serialization.cc:65-134).
// No device needed: prints the same "### NAME" blocks; tests/test_cpp_host_api.py compares text byte for byte and the JSON
// document both as text and as a parsed object.
#include <cstdio>

#include "mini_opt_b200/nonlinear.hpp"

using namespace mini_opt_b200;

static QPInteriorPointIteration QpIter(int i, bool probe) {
  double row[MO_QP_ITER_LOG_STRIDE] = {1e-9 * (i + 1), 0.1, 0.0, 2.5e16, 1e-3 / 3.0, 1e-5, 100000.0, 1e15,
                                       1.0 / (3.0 + i), 0.995 / (1 + i), 1.0,
                                       std::nan(""), std::nan(""), std::nan("")};
  if (probe) { row[11] = 0.25; row[12] = 1e-5; row[13] = 123456789.125; }
  return QPInteriorPointIteration::FromLogRow(row);  // the path a solve takes: one iter_log row
}

int main() {
  const QPInteriorPointIteration a = QpIter(0, false), b = QpIter(1, true);
  std::printf("### QP_ITER_PLAIN\n%s### QP_ITER_PROBE\n%s", a.ToString().c_str(), b.ToString().c_str());
  QPInteriorPointSolverOutputs ip_out;
  ip_out.termination_state = QPInteriorPointTerminationState::SATISFIED_KKT_TOL;
  ip_out.iterations = {a, b};
  ip_out.num_iterations = 2;
  ip_out.lagrange_multipliers = QPLagrangeMultipliers{-0.5, 2.0};
  QPInteriorPointSolverOutputs ip_out2;
  ip_out2.termination_state = QPInteriorPointTerminationState::MAX_ITERATIONS;
  ip_out2.iterations = {b};
  ip_out2.num_iterations = 1;
  NLSSolverOutputs out;
  out.termination_state = NLSTerminationState::SATISFIED_FIRST_ORDER_TOL;
  out.iterations.push_back(NLSIteration{0, OptimizerState::NOMINAL, 0.0, Errors{12.5, 0.3}, ip_out.termination_state, 2,
                                        DirectionalDerivatives{-3.25, -0.3}, 1.01, StepSizeSelectionResult::SUCCESS,
                                        {LineSearchStep{1.0, Errors{0.75, 1e-3}}, LineSearchStep{0.5, Errors{0.0625, 1e-12}}}, ip_out});
  out.iterations.push_back(NLSIteration{1, OptimizerState::ATTEMPTING_RESTORE_LM, 1e-3, Errors{0.0625, 1e-12},
                                        QPInteriorPointTerminationState::SATISFIED_KKT_TOL, 1, DirectionalDerivatives{-1e-7, 0.0},
                                        2.0 / 3.0, StepSizeSelectionResult::MAX_ITERATIONS, {}, QPNullSpaceTerminationState::SUCCESS});
  out.iterations.push_back(NLSIteration{2, OptimizerState::NOMINAL, 10.0, Errors{1e-20, 0.0}, ip_out2.termination_state, 1,
                                        DirectionalDerivatives{0.0, 0.0}, 1e6, StepSizeSelectionResult::FIRST_ORDER_SATISFIED,
                                        {LineSearchStep{0.125, Errors{1e-20, 0.0}}}, ip_out2});
  std::printf("### NLS_PLAIN\n%s", out.ToString(false, false).c_str());
  std::printf("### NLS_COLOR_QP\n%s", out.ToString(true, true).c_str());
  std::printf("### NLS_COUNTS\n%zu %zu %zu\n", out.NumQPIterations(), out.NumLineSearchSteps(), out.NumFailedLineSearches());
  std::printf("### NLS_JSON\n%s\n### END\n", out.ToJson().c_str());
  return 0;
}
