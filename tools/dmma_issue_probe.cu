// What limits DMMA.8x8x4 issue in a real 64x32 warp tile?  Register-only and LDS-fed variants of the
// engine's inner loop (8 warps/SM = 2 per sub-partition, 64 accumulator doubles per thread).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
// MODE 0: i outer / j inner, fragments fixed in registers
// MODE 1: j outer / i inner
// MODE 2: like 0, fragments re-loaded from shared memory every 64 DMMAs (12 LDS.128)
// MODE 3: like 2 with a 16x4 tile instead (128x... no) -- unused
template <int MODE>
__global__ void __launch_bounds__(256, 1) tile_loop(double* out, const double* in, int iters) {
  __shared__ __align__(16) double sm[4096];
  for (int i = threadIdx.x; i < 4096; i += 256) sm[i] = in[i];
  __syncthreads();
  double acc[8][4][2];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  double a[2][8], b[2][4];
  for (int e = 0; e < 2; ++e) { for (int i = 0; i < 8; ++i) a[e][i] = in[threadIdx.x + 256 * (e * 8 + i)]; for (int j = 0; j < 4; ++j) b[e][j] = in[threadIdx.x + 256 * (16 + e * 4 + j)]; }
  const double2* s2 = reinterpret_cast<const double2*>(sm) + (threadIdx.x & 31);
  for (int it = 0; it < iters; ++it) {
    if (MODE == 2) {
      const int o = (it & 7) * 32 * 12;
#pragma unroll
      for (int i = 0; i < 8; ++i) { double2 v = s2[(o + 32 * i) & 2047]; a[0][i] = v.x; a[1][i] = v.y; }
#pragma unroll
      for (int j = 0; j < 4; ++j) { double2 v = s2[(o + 32 * (8 + j)) & 2047]; b[0][j] = v.x; b[1][j] = v.y; }
    }
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      if (MODE == 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int i = 0; i < 8; ++i) dmma(acc[i][j], a[e][i], b[e][j]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma(acc[i][j], a[e][i], b[e][j]);
      }
    }
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += acc[i][j][0] + acc[i][j][1];
  if (s == 12345.678) out[0] = s;
}
template <int MODE>
static void run(const char* name, int sms, double* out, double* in) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000; float best = 1e30f;
  for (int r = 0; r < 4; ++r) {
    float ms; cudaEventRecord(e0); tile_loop<MODE><<<sms, 256>>>(out, in, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  printf("%-44s %.2f TFLOP/s (%s)\n", name, 512.0 * 64 * iters * 8 * sms / best * 1e-9, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  double *out, *in; cudaMalloc(&out, 8); cudaMalloc(&in, 8192 * 8); cudaMemset(in, 0, 8192 * 8);
  run<0>("regs, i outer / j inner (A reused x4)", p.multiProcessorCount, out, in);
  run<1>("regs, j outer / i inner (B reused x8)", p.multiProcessorCount, out, in);
  run<2>("i outer + 12 LDS.128 per 64 DMMA", p.multiProcessorCount, out, in);
  return 0;
}
