// Microbenchmark: FP64 pipe cost of a DFMA issued with 32 / 16 / 8 / 1 active lanes (divergent branch), and of a
// dependent DFMA chain running next to a DMMA stream of another warp on the same SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lanes_bench lanes_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ACTIVE>
__global__ void __launch_bounds__(128) dfma_lanes(double* out, int iters, double a, double b) {
  double r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x * 1e-3 + i;
  if ((threadIdx.x & 31) < ACTIVE) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = fma(r[i], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += r[i];
  if (s == 123456.789) out[0] = s;
}

// warp 0 of each CTA: dependent DFMA chain (latency probe); warps 1..: DMMA streams (NACC independent accumulators)
template <int NACC>
__global__ void __launch_bounds__(256) chain_vs_dmma(long long* cyc, double* out, int iters, double a, double b, int dmma_warps) {
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    double x = threadIdx.x;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int u = 0; u < 16; ++u) x = fma(x, a, b);
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (x == 123456.789) out[0] = x;
  } else if (warp <= dmma_warps) {
    double c[NACC][2];
#pragma unroll
    for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = i; }
    for (int it = 0; it < iters * 2; ++it) {
#pragma unroll
      for (int i = 0; i < NACC; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
    if (s == 123456.789) out[0] = s;
  }
}

template <class K>
float time_kernel(K launch) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep > 0 && ms < best) best = ms;
  }
  return best;
}

int main() {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* d; cudaMalloc(&d, 64); long long* cyc; cudaMalloc(&cyc, sizeof(long long) * sms * 8);
  const int iters = 4096, blocks = sms * 4, threads = 128;
  const double ops = 32.0 * iters * blocks * threads;  // DFMA thread-instructions if all lanes were active
  float ms;
  ms = time_kernel([&] { dfma_lanes<32><<<blocks, threads>>>(d, iters, 0.999, 1e-9); }); printf("DFMA active lanes 32: %.3f ms  (%.2f TFLOP/s counted on active lanes)\n", ms, 2 * ops / ms / 1e9);
  ms = time_kernel([&] { dfma_lanes<16><<<blocks, threads>>>(d, iters, 0.999, 1e-9); }); printf("DFMA active lanes 16: %.3f ms\n", ms);
  ms = time_kernel([&] { dfma_lanes<8><<<blocks, threads>>>(d, iters, 0.999, 1e-9); }); printf("DFMA active lanes  8: %.3f ms\n", ms);
  ms = time_kernel([&] { dfma_lanes<1><<<blocks, threads>>>(d, iters, 0.999, 1e-9); }); printf("DFMA active lanes  1: %.3f ms\n", ms);
  // dependent chain latency with 0 / 1 / 3 / 7 DMMA warps in the same CTA (one CTA per SM; warps 0,4 share SMSP 0)
  for (int dw : {0, 1, 3, 4, 7}) {
    chain_vs_dmma<4><<<sms, 256>>>(cyc, d, 2048, 0.999, 1e-9, dw);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("dependent DFMA latency with %d DMMA warps in the CTA: %.1f cycles\n", dw, (double)h / (2048.0 * 16));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
