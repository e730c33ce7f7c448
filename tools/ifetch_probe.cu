// Is a single warp running long straight-line code limited by instruction fetch?  Same work, once fully
// unrolled (STRAIGHT instructions in a row) and once as a loop over a short body.
#include <cstdio>
#include <cuda_runtime.h>
template <int UNROLL, int TRIPS>
__global__ void k(unsigned* out, long long* cyc, unsigned s) {
  unsigned a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = s + j + threadIdx.x;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < TRIPS; ++it) {
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = a[j] * 3u + (a[(j + 1) & 7] ^ (unsigned)i);   // IMAD + LOP3, 8 independent chains
    }
  }
  long long t1 = clock64();
  unsigned r = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) r ^= a[j];
  out[threadIdx.x] = r;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int U, int T> void run(const char* name, unsigned* out, long long* cyc) {
  for (int it = 0; it < 3; ++it) k<U, T><<<1, 32>>>(out, cyc, 1u);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-40s %8lld cycles for %d ops  -> %.2f cycles/op\n", name, h, U * T * 16, (double)h / (U * T * 16));
}
int main() {
  unsigned* out; long long* cyc; cudaMalloc(&out, 128); cudaMalloc(&cyc, 8);
  run<4, 128>("loop body 64 instr x 128 trips", out, cyc);
  run<16, 32>("loop body 256 instr x 32 trips", out, cyc);
  run<64, 8>("loop body 1024 instr x 8 trips", out, cyc);
  run<128, 4>("loop body 2048 instr x 4 trips", out, cyc);
  run<512, 1>("straight line 8192 instr", out, cyc);
  run<512, 4>("loop body 8192 instr x 4 trips", out, cyc);
  return cudaDeviceSynchronize() != cudaSuccess;
}
