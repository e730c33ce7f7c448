// Phase timing of the one-CTA potrf+trtri leaf (clock64 stamps) -- build and run on a B200:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -DCAPK_LEAF_PROF -o /tmp/leaf_probe tools/leaf_probe.cu && /tmp/leaf_probe
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#ifndef LEAF_HEADER
#define LEAF_HEADER "../capital_b200/csrc/potrf_leaf.cuh"
#endif
#include LEAF_HEADER

using namespace capk;

__global__ void evict_kernel(double* p, size_t n) {   // something else between leaves, as in the recursion
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = p[i] * 1.0000001 + 1.0;
}

int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 128;
  std::vector<double> A((size_t)n * n), R((size_t)n * n), X((size_t)n * n);
  srand(1);
  for (int c = 0; c < n; ++c)
    for (int r = 0; r <= c; ++r) {
      double v = (rand() / (double)RAND_MAX - 0.5);
      if (r == c) v = n + fabs(v);
      A[r + (size_t)c * n] = A[c + (size_t)r * n] = v;
    }
  double *dA, *dW, *dX, *dE;
  cudaMalloc(&dA, A.size() * 8); cudaMalloc(&dW, A.size() * 8); cudaMalloc(&dX, A.size() * 8);
  const size_t ne = 64u << 20; cudaMalloc(&dE, ne * 8); cudaMemset(dE, 0, ne * 8);
  int* dinfo; cudaMalloc(&dinfo, 4); cudaMemset(dinfo, 0, 4);
  cudaMemcpy(dA, A.data(), A.size() * 8, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(potrf_trtri_base_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BASE_SMEM_BYTES);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode) {          // 0: back-to-back leaves, 1: another kernel in between
    for (int it = 0; it < 6; ++it) {
      cudaMemcpy(dW, dA, A.size() * 8, cudaMemcpyDeviceToDevice);
      if (mode) evict_kernel<<<1184, 256>>>(dE, ne);
      cudaEventRecord(e0);
      potrf_trtri_base_kernel<<<1, BASE_THREADS, BASE_SMEM_BYTES>>>(dW, n, dX, n, n, dinfo, 0, 1, 1, 0, 0);
      cudaEventRecord(e1);
      cudaError_t err = cudaDeviceSynchronize();
      if (err != cudaSuccess) { printf("error %s\n", cudaGetErrorString(err)); return 1; }
      float ms; cudaEventElapsedTime(&ms, e0, e1);
#ifdef CAPK_LEAF_PROF
      long long st[64]; cudaMemcpyFromSymbol(st, g_leaf_prof, sizeof(st));
      const int nb = (n + PB - 1) / PB;
      printf("mode %d it %d: %.1f us | load %lld |", mode, it, ms * 1e3, st[1] - st[0]);
      for (int b = 0; b < nb; ++b) printf(" b%d %lld+%lld", b, st[2 + 2 * b] - st[1 + 2 * b], st[3 + 2 * b] - st[2 + 2 * b]);
      printf(" | store %lld | total %lld cycles\n", st[30] - st[1 + 2 * nb], st[30] - st[0]);
#ifdef CAPK_LEAF_REPEAT
      printf("    S2(0) repeated: %lld %lld %lld   sweep(0) repeated: %lld %lld %lld\n", st[40] - st[2], st[41] - st[40], st[42] - st[41], st[44] - st[1], st[45] - st[44], st[46] - st[45]);
#endif
#else
      printf("mode %d it %d: %.1f us\n", mode, it, ms * 1e3);
#endif
    }
  }
  // residual check
  cudaMemcpy(R.data(), dW, R.size() * 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(X.data(), dX, X.size() * 8, cudaMemcpyDeviceToHost);
  double e_chol = 0, e_inv = 0;
  for (int c = 0; c < n; ++c)
    for (int r = 0; r <= c; ++r) {
      double s = 0, q = 0;
      for (int k = 0; k <= r; ++k) s += R[k + (size_t)r * n] * R[k + (size_t)c * n];
      for (int k = r; k <= c; ++k) q += R[r + (size_t)k * n] * X[k + (size_t)c * n];
      e_chol = fmax(e_chol, fabs(s - A[r + (size_t)c * n]));
      e_inv = fmax(e_inv, fabs(q - (r == c ? 1.0 : 0.0)));
    }
  int info; cudaMemcpy(&info, dinfo, 4, cudaMemcpyDeviceToHost);
  printf("n=%d max|R^T R - A| = %.3e  max|R X - I| = %.3e  info %d\n", n, e_chol, e_inv, info);
  return 0;
}
