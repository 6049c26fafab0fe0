// Microbenchmark: FP64 throughput of DFMA vs DMMA (mma.sync m8n8k4 / m16n8k8 f64) on sm_100a, and LDS-fed variants.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_peak dmma_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double a, double b) {
  double r[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) r[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int i = 0; i < 16; ++i) r[i] = fma(r[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += r[i];
  if (s == 123456.789) out[0] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) dmma884_kernel(double* out, int iters, double a, double b) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
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

template <int NACC>
__global__ void __launch_bounds__(256) dmma1688_kernel(double* out, int iters, double a, double b) {
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123456.789) out[0] = s;
}

// dependent chain latency of DMMA and DFMA (1 warp, 1 accumulator)
__global__ void lat_kernel(long long* out, double a, double b) {
  double c0 = threadIdx.x, c1 = 1.0;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < 1024; ++it)
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
  long long t1 = clock64();
  double r = c0;
#pragma unroll 1
  for (int it = 0; it < 1024; ++it) r = fma(r, a, b);
  long long t2 = clock64();
  double q = r;
#pragma unroll 1
  for (int it = 0; it < 256; ++it) q = 1.0 / q + a;
  long long t3 = clock64();
  if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t1; out[2] = t3 - t2; out[3] = (long long)(c0 + c1 + r + q); }
}

template <typename F>
double time_ms(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 1e30;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r > 0 && ms < best) best = ms;
  }
  return best;
}

int main() {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* d; cudaMalloc(&d, 64);
  const int iters = 4096;
  for (int bps : {1, 2, 4, 8}) {
    const int blocks = sms * bps, threads = 256;
    double ms = time_ms([&] { dfma_kernel<<<blocks, threads>>>(d, iters, 0.999999, 1e-9); });
    printf("DFMA      blocks/SM=%d: %.2f TFLOP/s\n", bps, 2.0 * 32 * iters * blocks * threads / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { dmma884_kernel<4><<<blocks, threads>>>(d, iters, 0.999999, 1e-9); });
    printf("DMMA884x4 blocks/SM=%d: %.2f TFLOP/s\n", bps, 512.0 * 4 * iters * blocks * (threads / 32) / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { dmma884_kernel<8><<<blocks, threads>>>(d, iters, 0.999999, 1e-9); });
    printf("DMMA884x8 blocks/SM=%d: %.2f TFLOP/s\n", bps, 512.0 * 8 * iters * blocks * (threads / 32) / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { dmma1688_kernel<4><<<blocks, threads>>>(d, iters, 0.999999, 1e-9); });
    printf("DMMA1688x4 blocks/SM=%d: %.2f TFLOP/s\n", bps, 2048.0 * 4 * iters * blocks * (threads / 32) / (ms * 1e-3) / 1e12);
  }
  // one warp per SMSP only
  {
    double ms = time_ms([&] { dmma884_kernel<8><<<sms, 128>>>(d, iters, 0.999999, 1e-9); });
    printf("DMMA884x8 1 warp/SMSP: %.2f TFLOP/s\n", 512.0 * 8 * iters * sms * 4 / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { dmma884_kernel<1><<<sms, 128>>>(d, iters, 0.999999, 1e-9); });
    printf("DMMA884x1 1 warp/SMSP: %.2f TFLOP/s\n", 512.0 * 1 * iters * sms * 4 / (ms * 1e-3) / 1e12);
  }
  long long* l; cudaMalloc(&l, 64); long long h[4];
  lat_kernel<<<1, 32>>>(l, 0.999999, 1e-9); cudaMemcpy(h, l, 32, cudaMemcpyDeviceToHost);
  printf("latency: DMMA884 dependent %.1f cycles, DFMA dependent %.1f cycles, (1/x + a) %.1f cycles\n", h[0] / 1024.0, h[1] / 1024.0, h[2] / 256.0);
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
