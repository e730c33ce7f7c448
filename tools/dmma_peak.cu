// FP64 peak calibration for the roofline: DMMA.8x8x4 issue-rate and DFMA-rate microkernels (no memory traffic).
// Prints TFLOP/s; run alone on one GPU.  MEASURED_PEAKS.json has no FP64 entry, so this is its source.
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void dmma_loop(double* out, int iters, double seed) {
  double c[ILP][2]; double a = seed + threadIdx.x, b = seed - threadIdx.x;
  for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = 0.0;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  double s = 0; for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
  if (s == 12345.678) out[0] = s;
}
template <int ILP>
__global__ void dfma_loop(double* out, int iters, double seed) {
  double c[ILP]; double a = seed + threadIdx.x * 1e-9, b = seed * 0.5;
  for (int i = 0; i < ILP; ++i) c[i] = i;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < ILP; ++i) c[i] = fma(c[i], a, b);
  double s = 0; for (int i = 0; i < ILP; ++i) s += c[i];
  if (s == 12345.678) out[0] = s;
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double* out; cudaMalloc(&out, 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int warps = 4; warps <= 16; warps *= 2) {
    float best_m = 1e30f, best_f = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
      float ms;
      cudaEventRecord(e0); dmma_loop<16><<<p.multiProcessorCount, warps * 32>>>(out, iters, 1.0); cudaEventRecord(e1); cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1); if (ms < best_m) best_m = ms;
      cudaEventRecord(e0); dfma_loop<16><<<p.multiProcessorCount, warps * 32>>>(out, iters * 8, 1.0); cudaEventRecord(e1); cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1); if (ms < best_f) best_f = ms;
    }
    double fl_m = 2.0 * 256 * 16 * (double)iters * warps * p.multiProcessorCount;
    double fl_f = 2.0 * 32 * 16 * (double)iters * 8 * warps * p.multiProcessorCount;
    printf("warps/SM=%2d  DMMA %.2f TFLOP/s   DFMA %.2f TFLOP/s\n", warps, fl_m / best_m * 1e-9, fl_f / best_f * 1e-9);
  }
  // sustained 3 s DMMA run (power-capped clock)
  cudaEventRecord(e0);
  int launches = 0; float ms = 0;
  do { dmma_loop<16><<<p.multiProcessorCount, 512>>>(out, iters * 10, 1.0); ++launches; cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1); } while (ms < 3000.f);
  printf("sustained DMMA (%.1f s, 16 warps/SM): %.2f TFLOP/s\n", ms * 1e-3, 2.0 * 256 * 16 * (double)iters * 10 * 16 * p.multiProcessorCount * launches / ms * 1e-9);
  printf("device=%s sms=%d clock_khz=%d err=%s\n", p.name, p.multiProcessorCount, p.clockRate, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
