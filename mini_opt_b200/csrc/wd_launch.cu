// Launcher of the warp-per-problem DMMA kernel (qp_wdm.cuh) in its own translation unit: the kernel is the one under
// active tuning and this keeps its rebuild at seconds. Called from dispatch_solve in capi.cu.
#include <cstdlib>

#include "qp_wdm.cuh"
#include "wd_launch.h"

namespace mo {

// (N, K, warps per problem, min CTAs per SM); MO_QP_WD_WARPS=1|2 overrides the warp count for A/B runs.
// Measured on B200 (ms per batch): n=64 64K QPs: 6 CTAs/SM (smem limit) 18.9; n=32 256K: 12 CTAs/SM (168 registers, no
// spills) 18.5, 16 (128 registers) 19.6, 20 (96) 23.4; n=16 512K: 12 -> 19.3, 16 -> 19.1, 24 -> 23.5.
#ifndef MO_WD_LA64
#define MO_WD_LA64 false
#endif
#ifndef MO_WD_LA32
#define MO_WD_LA32 false
#endif
#ifndef MO_WD_LA16
#define MO_WD_LA16 false
#endif
#ifndef MO_WD_PBS
#define MO_WD_PBS true
#endif
#define MO_WD_SHAPES(X) X(64, 16, 1, 7, MO_WD_LA64) X(32, 8, 1, 12, MO_WD_LA32) X(16, 8, 1, 16, MO_WD_LA16)
#ifdef MO_WD_AB_BUILD
#define MO_WD_ALT_SHAPES(X) X(64, 16, 2, 6, false) X(32, 8, 2, 12, false) X(16, 8, 2, 12, false)
#else
#define MO_WD_ALT_SHAPES(X)
#endif

template <int N, int K, int NW, int MINB, bool LA>
static cudaError_t launch_wd(const QPArgs& a, int sms, cudaStream_t stream) {
  using L = WdLayout<N, K>;
  auto kernel = qp_wd_kernel<N, K, NW, MINB, LA, (NW == 1) && MO_WD_PBS>;
  static thread_local int configured_dev = -1;
  static thread_local int bps = 0;
  int cur = 0;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return e;
  if (configured_dev != cur) {
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 32 * NW, L::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    if (bps < 1) return cudaErrorInvalidConfiguration;
    configured_dev = cur;
  }
  const long long cap = (long long)sms * bps;
  kernel<<<(int)(a.batch < cap ? a.batch : cap), 32 * NW, L::SMEM_BYTES, stream>>>(a);
  return cudaGetLastError();
}

bool wd_has_shape(int n, int k) {
#define X(N_, K_, W_, B_, LA_) if (n == N_ && k == K_) return true;
  MO_WD_SHAPES(X)
#undef X
  return false;
}

cudaError_t wd_launch(const QPArgs& a, int sms, cudaStream_t stream) {
#ifdef MO_WD_AB_BUILD  // lab build only: MO_QP_WD_WARPS=1|2 selects the one- / two-warp variant (the product build has one kernel per shape)
  const char* e = getenv("MO_QP_WD_WARPS");
  const int want = e ? atoi(e) : 0;
#else
  constexpr int want = 0;
#endif
#define X(N_, K_, W_, B_, LA_) if (a.n == N_ && a.k == K_ && (want == 0 || want == W_)) return launch_wd<N_, K_, W_, B_, LA_>(a, sms, stream);
  MO_WD_SHAPES(X)
  MO_WD_ALT_SHAPES(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace mo
