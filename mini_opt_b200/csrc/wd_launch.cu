// Launcher of the warp-per-problem DMMA kernel (qp_wdm.cuh) in its own translation unit: the kernel is the one under
// active tuning and this keeps its rebuild at seconds. Called from dispatch_solve in capi.cu.
#include <atomic>
#include <cstdlib>

#include "qp_wdm.cuh"
#include "wd_launch.h"

namespace mo {

// Residency, measured on B200 in round 1 (ms per batch): n=64 64K QPs: 6 CTAs/SM 18.9, 7 (shared memory limit) 18.4;
// n=32 256K: 12 CTAs/SM (168 registers, no spills) 18.5, 16 (128 registers) 19.6, 20 (96) 23.4; n=16 512K: 12 -> 19.3,
// 16 -> 19.1, 24 -> 23.5.
// Schedules, measured on B200 in round 2 (profiles/r02_wd_variants.txt, ms per batch of 64K / 256K / 512K QPs):
//   plain left-looking columns, runtime back substitution loop      n=64 18.65   n=32 18.16   n=16 18.95   <- 64, 32
//   + software-pipelined, unrolled back substitution (PBS)          n=64 19.23   n=32 18.37   n=16 18.43   <- 16
//   + look-ahead columns (LA: pivot chain under the gather DMMAs)    n=64 21.38   n=32 18.37   n=16 18.41
// LA and PBS lose at n = 64 on instruction fetch: the kernel grows from 9.1K to 12.4K SASS instructions and
// smsp stall no_instruction doubles (0.55 -> 1.12 cycles per issue, profiles/r02_wd_variants.txt).
// (N, K, warps per problem, min CTAs per SM, look-ahead columns, pipelined back substitution)
#ifndef MO_WD_SHAPES
#define MO_WD_SHAPES(X) X(64, 16, 1, 7, false, false) X(32, 8, 1, 12, false, false) X(16, 8, 1, 16, false, true)
#endif
#ifdef MO_WD_AB_BUILD
#define MO_WD_ALT_SHAPES(X) X(64, 16, 2, 6, false, false) X(32, 8, 2, 12, false, false) X(16, 8, 2, 12, false, false)
#else
#define MO_WD_ALT_SHAPES(X)
#endif

#ifndef MO_WD_LOCK
#define MO_WD_LOCK 0
#endif
static std::atomic<int> g_wd_lock{MO_WD_LOCK};  // mo_debug_set_wd_lock: 1 = lock-step CTAs of MINB warps (measured SLOWER, kept for the A/B), 0 = one-warp CTAs

// Lock-step variant: ONE CTA of MINB warps per SM (the residency of the one-warp kernel), a problem per warp.
template <int N, int K, int MINB, bool PBS>
static cudaError_t launch_wd_lock(const QPArgs& a, int sms, cudaStream_t stream) {
  auto kernel = qp_wd_kernel<N, K, 1, 1, false, PBS, MINB>;
  constexpr size_t smem = sizeof(double) * wd_region_doubles<N, K>() * MINB;
  static thread_local int configured_dev = -1;
  int cur = 0;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return e;
  if (configured_dev != cur) {
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured_dev = cur;
  }
  const long long want = (a.batch + MINB - 1) / MINB;
  kernel<<<(int)(want < sms ? want : sms), 32 * MINB, smem, stream>>>(a);
  return cudaGetLastError();
}

template <int N, int K, int NW, int MINB, bool LA, bool PBS>
static cudaError_t launch_wd(const QPArgs& a, int sms, cudaStream_t stream) {
  using L = WdLayout<N, K>;
  if constexpr (NW == 1 && !LA) {
    if (g_wd_lock.load()) return launch_wd_lock<N, K, MINB, PBS>(a, sms, stream);
  }
  auto kernel = qp_wd_kernel<N, K, NW, MINB, LA, PBS>;
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
#define X(N_, K_, W_, B_, LA_, PBS_) if (n == N_ && k == K_) return true;
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
#define X(N_, K_, W_, B_, LA_, PBS_) if (a.n == N_ && a.k == K_ && (want == 0 || want == W_)) return launch_wd<N_, K_, W_, B_, LA_, PBS_>(a, sms, stream);
  MO_WD_SHAPES(X)
  MO_WD_ALT_SHAPES(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace mo

// Measurement hook: lock-step CTAs of the warp-per-problem DMMA kernel (1) or one-warp CTAs (0, default). Results are
// bit-identical; lock-step measured slower (profiles/r02_wd_lockstep_ab.txt). Returns the previous value.
extern "C" int mo_debug_set_wd_lock(int enabled) { return mo::g_wd_lock.exchange(enabled ? 1 : 0); }
