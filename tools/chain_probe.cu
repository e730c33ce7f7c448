// Latency of the pieces of the pivot chain, single warp, every op an asm volatile so that nothing moves
// across the clock reads.
#include <cstdio>
#include <cuda_runtime.h>
#define CLK(t) asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory")
__device__ __forceinline__ double vfma(double a, double b, double c) { double r; asm volatile("fma.rn.f64 %0, %1, %2, %3;" : "=d"(r) : "d"(a), "d"(b), "d"(c)); return r; }
__device__ __forceinline__ double vmul(double a, double b) { double r; asm volatile("mul.rn.f64 %0, %1, %2;" : "=d"(r) : "d"(a), "d"(b)); return r; }
__device__ __forceinline__ double vrcp(double a) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a)); return r; }
__device__ __forceinline__ double vshfl(double a, int src) {
  int lo = __double2loint(a), hi = __double2hiint(a), lo2, hi2;
  asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(lo2) : "r"(lo), "r"(src));
  asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(hi2) : "r"(hi), "r"(src));
  return __hiloint2double(hi2, lo2);
}
constexpr int N = 128;
__global__ void probe(double* out, long long* cyc, double x0) {
  __shared__ double sh[64];
  const int lane = threadIdx.x;
  double x = x0 + lane * 1e-3;
  long long t0, t1;
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) x = vfma(x, 1.0000001, 1e-9);
  CLK(t1); if (lane == 0) cyc[0] = t1 - t0;
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) x = vmul(x, 1.0000001);
  CLK(t1); if (lane == 0) cyc[1] = t1 - t0;
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) x = vrcp(x);
  CLK(t1); if (lane == 0) cyc[2] = t1 - t0;
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) x = vshfl(x, (lane + 1) & 31);
  CLK(t1); if (lane == 0) cyc[3] = t1 - t0;
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) {
    asm volatile("st.shared.f64 [%0], %1;" :: "r"((unsigned)__cvta_generic_to_shared(&sh[lane + (i & 1) * 32])), "d"(x) : "memory");
    __syncwarp();
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"((unsigned)__cvta_generic_to_shared(&sh[((i * 7) & 31) + (i & 1) * 32])) : "memory");
  }
  CLK(t1); if (lane == 0) cyc[4] = t1 - t0;
  // the pivot chain: shfl -> rcp seed -> 3 fma -> fma
  x = 1.5 + lane * 1e-3;
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double d = vshfl(x, (i + 1) & 31);
    double c0 = vrcp(d);
    double e = vfma(-d, c0, 1.0);
    double e2 = vfma(e, e, e);
    double c = vfma(c0, e2, c0);
    x = vfma(-0.25, c, x + 1.0);
  }
  CLK(t1); if (lane == 0) cyc[5] = t1 - t0;
  // same without the shuffle
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double d = x;
    double c0 = vrcp(d);
    double e = vfma(-d, c0, 1.0);
    double e2 = vfma(e, e, e);
    double c = vfma(c0, e2, c0);
    x = vfma(-0.25, c, 2.0);
  }
  CLK(t1); if (lane == 0) cyc[6] = t1 - t0;
  // dependent DSETP + FSEL pair
  CLK(t0);
#pragma unroll
  for (int i = 0; i < N; ++i) { x = (x > 1.0) ? vmul(x, 0.999) : x; }
  CLK(t1); if (lane == 0) cyc[7] = t1 - t0;
  out[lane] = x;
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 16 * 8);
  for (int it = 0; it < 2; ++it) probe<<<1, 32>>>(out, cyc, 1.5);
  long long h[16]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[] = {"DFMA", "DMUL", "MUFU.RCP64H (+moves)", "SHFL f64", "STS->syncwarp->LDS", "shfl+rcp+3fma+fma", "rcp+3fma+fma", "dsetp+predicated dmul"};
  for (int i = 0; i < 8; ++i) printf("%-28s %7.1f cycles\n", names[i], (double)h[i] / N);
  return cudaDeviceSynchronize() != cudaSuccess;
}
