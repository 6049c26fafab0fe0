// 1-D bulk asynchronous copies (TMA, cp.async.bulk) + mbarrier completion for staging the contiguous per-problem
// arrays of a batch (G, c, A_eq, b_eq, bounds / constraint lists; dense Jacobians) into shared memory.
//
// These replace, per problem, the element-by-element reads the reference does through Eigen at
// source/qp.cc:289-292 (G lower triangle, A_eq into H) and :404-409 (G, c, A_eq, b_eq in the KKT residual): one
// instruction per array moves the whole slice, the data lands without passing through registers, and the mbarrier lets
// the copy of problem p+1 run under the solve of problem p. SASS: UBLKCP + SYNCS (no LDGSTS / LDG in the load path).
//
// Constraints of cp.async.bulk: source, destination and size are multiples of 16 bytes. A slice that is not (odd
// element counts, unaligned base) is loaded by the lanes with plain loads into the same buffer (`stage_array`).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mo {

__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the initialised barrier visible to the async proxy (the TMA unit)
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  while (!mbar_try_wait(bar, parity)) {}
}

// generic-proxy accesses (ld/st.shared) of this thread are ordered before later async-proxy (TMA) accesses
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// global -> shared bulk copy; completion (bytes) is signalled on `bar`
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ bool bulk_ok(const void* gsrc, size_t bytes) {
  return bytes > 0 && (bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(gsrc) & 15u) == 0;
}

// Stage one contiguous slice of `bytes` bytes (a multiple of 4) into `dst` (16-byte aligned). With `tma` and an
// eligible slice lane 0 issues one bulk copy and returns the byte count to add to the barrier's expected transaction
// (the caller sums these and arms the barrier BEFORE... after issuing is also legal: the phase cannot complete before
// the pending arrival of expect_tx); otherwise all 32 lanes copy it with plain loads and 0 is returned.
// Must be called by the whole warp.
__device__ __forceinline__ unsigned stage_array(void* dst, const void* gsrc, size_t bytes, uint64_t* bar, bool tma, int lane) {
  if (bytes == 0 || gsrc == nullptr) return 0u;
  if (tma && bulk_ok(gsrc, bytes)) {
    if (lane == 0) bulk_g2s(dst, gsrc, static_cast<unsigned>(bytes), bar);
    return static_cast<unsigned>(bytes);
  }
  const unsigned words = static_cast<unsigned>(bytes >> 2);
  const uint32_t* __restrict__ s = static_cast<const uint32_t*>(gsrc);
  uint32_t* __restrict__ d = static_cast<uint32_t*>(dst);
  if ((bytes & 7u) == 0 && (reinterpret_cast<uintptr_t>(gsrc) & 7u) == 0) {
    const double* __restrict__ s8 = static_cast<const double*>(gsrc);
    double* __restrict__ d8 = static_cast<double*>(dst);
    for (unsigned i = lane; i < (words >> 1); i += 32) d8[i] = __ldg(s8 + i);
  } else {
    for (unsigned i = lane; i < words; i += 32) d[i] = __ldg(s + i);
  }
  return 0u;
}

__host__ __device__ constexpr size_t align16(size_t v) { return (v + 15) & ~size_t(15); }

}  // namespace mo
