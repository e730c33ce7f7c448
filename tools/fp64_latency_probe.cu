// Single-warp latencies that bound the serial part of the leaf Cholesky: dependent DFMA / DMUL / MUFU seeds /
// SHFL / shared-memory round trip, the issue rate of independent DFMAs from one warp, and the accuracy of the
// MUFU.RCP64H / MUFU.RSQ64H seeds.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/fp64_latency_probe tools/fp64_latency_probe.cu
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>

__device__ __forceinline__ double rcp_seed(double d) { double y; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d)); return y; }
__device__ __forceinline__ double rsq_seed(double d) { double y; asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d)); return y; }

constexpr int N = 256;

__global__ void probe(double* out, long long* cyc, double x0) {
  __shared__ double sh[64];
  const int lane = threadIdx.x;
  double x = x0 + lane * 1e-9, acc = 0.0;
  long long t0, t1;
  // 0: dependent DFMA
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) x = fma(x, 1.0000001, 1e-9);
  t1 = clock64(); if (lane == 0) cyc[0] = t1 - t0; acc += x;
  // 1: dependent DMUL
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) x = x * 1.0000001;
  t1 = clock64(); if (lane == 0) cyc[1] = t1 - t0; acc += x;
  // 2: dependent rcp seed
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) x = rcp_seed(x);
  t1 = clock64(); if (lane == 0) cyc[2] = t1 - t0; acc += x;
  // 3: dependent rsqrt seed
  x = fabs(x) + 1.0;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) x = rsq_seed(x);
  t1 = clock64(); if (lane == 0) cyc[3] = t1 - t0; acc += x;
  // 4: dependent shuffle of a double
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) x = __shfl_sync(0xffffffffu, x, (i * 7) & 31);
  t1 = clock64(); if (lane == 0) cyc[4] = t1 - t0; acc += x;
  // 5: shared round trip STS -> syncwarp -> LDS (broadcast)
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) { sh[lane + (i & 1) * 32] = x; __syncwarp(); x = sh[((i * 7) & 31) + (i & 1) * 32] + 0.0; }
  t1 = clock64(); if (lane == 0) cyc[5] = t1 - t0; acc += x;
  // 6: 8 independent DFMA chains (issue rate)
  double z[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) z[j] = x + j;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) z[j] = fma(z[j], 1.0000001, 1e-9);
  t1 = clock64(); if (lane == 0) cyc[6] = t1 - t0;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc += z[j];
  // 7: 16 independent DFMA chains
  double w[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) w[j] = x + j;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < 16; ++j) w[j] = fma(w[j], 1.0000001, 1e-9);
  t1 = clock64(); if (lane == 0) cyc[7] = t1 - t0;
#pragma unroll
  for (int j = 0; j < 16; ++j) acc += w[j];
  // 8: dependent DFMA whose other operand comes from a broadcast LDS issued in the same iteration
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < N; ++i) x = fma(x, sh[i & 63], 1e-9);
  t1 = clock64(); if (lane == 0) cyc[8] = t1 - t0; acc += x;
  out[lane] = acc;
}

__global__ void seeds(double* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const double d = 1.0 + (i + 0.5) / (gridDim.x * blockDim.x);          // [1, 2)
  const double d4 = 4.0 * d;
  double e1 = fabs(rcp_seed(d) * d - 1.0), e2 = fabs(rsq_seed(d) * rsq_seed(d) * d - 1.0), e3 = fabs(rsq_seed(d4) * rsq_seed(d4) * d4 - 1.0);
  for (int o = 16; o; o >>= 1) {
    e1 = fmax(e1, __shfl_xor_sync(0xffffffffu, e1, o));
    e2 = fmax(e2, __shfl_xor_sync(0xffffffffu, e2, o));
    e3 = fmax(e3, __shfl_xor_sync(0xffffffffu, e3, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax((unsigned long long*)&err[0], (unsigned long long)__double_as_longlong(e1));
    atomicMax((unsigned long long*)&err[1], (unsigned long long)__double_as_longlong(e2));
    atomicMax((unsigned long long*)&err[2], (unsigned long long)__double_as_longlong(e3));
  }
}

int main() {
  double* out; long long* cyc; double* err;
  cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 16 * 8); cudaMalloc(&err, 3 * 8); cudaMemset(err, 0, 24);
  for (int it = 0; it < 2; ++it) probe<<<1, 32>>>(out, cyc, 1.5);
  seeds<<<4096, 256>>>(err);
  long long h[16]; double e[3];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost); cudaMemcpy(e, err, sizeof(e), cudaMemcpyDeviceToHost);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
  const char* names[] = {"dependent DFMA", "dependent DMUL", "dependent MUFU.RCP64H", "dependent MUFU.RSQ64H", "dependent SHFL (f64 = 2 x b32)",
                         "STS->syncwarp->LDS->DADD", "8 independent DFMA (per instr)", "16 independent DFMA (per instr)", "DFMA with LDS operand"};
  const int per[] = {N, N, N, N, N, N, 8 * N, 16 * N, N};
  for (int i = 0; i < 9; ++i) printf("%-34s %7.1f cycles\n", names[i], (double)h[i] / per[i]);
  printf("seed rel. error: rcp %.3e (2^%.1f)  rsqrt^2 %.3e (2^%.1f) / %.3e\n", e[0], log2(e[0]), e[1], log2(e[1]), e[2]);
  return 0;
}
