// Microbenchmark: cycles of one wd_diag_block call (LDL' of an 8x8 diagonal tile + inverse of L, qp_wdm.cuh) and of one
// wd_forward_block, in isolation: 1 .. 8 single-warp CTAs per SM, each looping over its own tile in shared memory.
// Answers whether the 2.24K cycles per block measured inside the kernel (profiles/r02_wd64_phase_trace.txt) are intrinsic to
// the routine or come from sharing the SM with other warps' DMMA phases.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../mini_opt_b200/csrc -I../../include -o diag_block_bench diag_block_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

#include "qp_wdm.cuh"

using namespace mo;

__global__ void __launch_bounds__(32) diag_bench(long long* cyc, double* sink, int iters) {
  extern __shared__ __align__(16) double sm[];
  using L = WdLayout<64, 16>;
  const int lane = threadIdx.x;
  double* tile = sm + L::tile(0, 0);
  double* ref = sm + L::tile(1, 0);  // pristine copy
  for (int e = lane; e < 64; e += 32) {
    const int r = e >> 3, c = e & 7;
    ref[wd_el(r, c)] = (r == c ? 10.0 + r : 0.3 / (1.0 + r + c));
  }
  for (int e = lane; e < L::P; e += 32) { sm[L::O_R + e] = 1.0 + e; sm[L::O_W + e] = 0.0; }
  __syncwarp();
  double chk = 0.0;
  long long t_diag = 0, t_fwd = 0;
  for (int it = 0; it < iters; ++it) {
    tile[lane] = ref[lane]; tile[32 + lane] = ref[32 + lane];
    __syncwarp();
    const long long t0 = clock64();
    wd_diag_block(tile, sm + L::O_D, sm + L::O_INVD, chk);
    const long long t1 = clock64();
    wd_forward_block<64, 16, 0>(sm, lane);
    const long long t2 = clock64();
    t_diag += t1 - t0; t_fwd += t2 - t1;
  }
  if (lane == 0) { cyc[2 * blockIdx.x] = t_diag; cyc[2 * blockIdx.x + 1] = t_fwd; }
  if (chk != chk) sink[0] = chk;
  sink[1 + blockIdx.x] = sm[L::O_W] + tile[wd_el(7, 0)];
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const size_t smem = WdLayout<64, 16>::SMEM_BYTES;
  cudaFuncSetAttribute(diag_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  long long* cyc;
  double* sink;
  const int iters = 2000;
  cudaMalloc(&cyc, sizeof(long long) * 2 * sms * 8);
  cudaMalloc(&sink, sizeof(double) * (1 + sms * 8));
  for (int per_sm = 1; per_sm <= 7; ++per_sm) {
    const int grid = sms * per_sm;
    diag_bench<<<grid, 32, smem>>>(cyc, sink, iters);
    cudaDeviceSynchronize();
    diag_bench<<<grid, 32, smem>>>(cyc, sink, iters);
    if (cudaDeviceSynchronize() != cudaSuccess) { std::printf("launch failed\n"); return 1; }
    long long* h = new long long[2 * grid];
    cudaMemcpy(h, cyc, sizeof(long long) * 2 * grid, cudaMemcpyDeviceToHost);
    double d = 0, f = 0;
    for (int i = 0; i < grid; ++i) { d += h[2 * i]; f += h[2 * i + 1]; }
    std::printf("warps per SM %d: wd_diag_block %.0f cycles, wd_forward_block %.0f cycles per call\n", per_sm, d / grid / iters, f / grid / iters);
    delete[] h;
  }
  return 0;
}
