// Interface between capi.cu and wd_launch.cu (warp-per-problem DMMA kernel).
#pragma once
#include "qp_common.cuh"

namespace mo {
bool wd_has_shape(int n, int k);                                       // an instantiation exists
cudaError_t wd_launch(const QPArgs& a, int sms, cudaStream_t stream);  // one kernel launch
}  // namespace mo
