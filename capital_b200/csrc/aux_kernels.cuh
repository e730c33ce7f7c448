// Small-block Cholesky + triangular inverse (the recursion's base case) and the HBM-bound helper
// kernels of the path (pack/unpack, copies, masks, axpby, generators).
//   potrf + trtri base case   lapack::engine::_potrf/_trtri  (reference src/lapack/interface.hpp:31-58,
//                             called together at cacqr.hpp:20-22 and cholinv/policy.h:199-201)
//   pack/unpack/copy          serialize<S1,S2>::invoke        (src/matrix/serialize.hpp:12-150)
//   mask                      util::remove_triangle           (src/util/util.hpp:267-293)
//   axpby                     summa.hpp:33,153
//   generators                rect::_distribute_random/_symmetric (src/matrix/structure.hpp:69-129)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace capk {

constexpr int BASE_NB = 128;                 // largest block factored by one CTA
constexpr int BASE_LD = BASE_NB + 1;
constexpr int BASE_THREADS = 512;
constexpr int BASE_SMEM_BYTES = (BASE_NB * BASE_LD + BASE_NB) * 8;

// One CTA: R = chol(A) (upper, A = R^T R) and X = R^{-1} for an n x n block, n <= 128.
// Shared buffer S: upper triangle holds A then R; the strict lower triangle holds X^T; xd = diag(X).
// Reads only the upper triangle of A.  Writes R into the upper triangle of A (strict lower untouched,
// like LAPACK) and X into the upper triangle of Rinv (strict lower zero-filled when zero_lower).
// info (device int, may be null): first non-positive pivot, 1-based, offset by info_base; 0 = ok.
__global__ void __launch_bounds__(BASE_THREADS, 1)
potrf_trtri_base_kernel(double* A, int64_t lda, double* Rinv, int64_t ldri, int n, int* info, int info_base,
                        int write_r, int zero_lower, int skip_potrf, int lower) {
  // lower != 0: the block is lower triangular; it is read and written transposed, so the same
  // upper-triangular code yields L = chol(A) (A = L L^T) and L^{-1}.
  extern __shared__ __align__(16) double sh[];
  double* S = sh;
  double* xd = sh + BASE_NB * BASE_LD;
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  if (tid == 0) s_bad = 0;
  for (int idx = tid; idx < n * n; idx += BASE_THREADS) {
    const int r = idx % n, c = idx / n;
    S[r * BASE_LD + c] = (r <= c) ? (lower ? A[c + (int64_t)r * lda] : A[r + (int64_t)c * lda]) : 0.0;
  }
  __syncthreads();
  // right-looking Cholesky on the upper triangle
  for (int j = 0; j < (skip_potrf ? 0 : n); ++j) {
    const double d = S[j * BASE_LD + j];
    if (!(d > 0.0)) { if (tid == 0 && s_bad == 0) s_bad = j + 1; }
    const double rd = 1.0 / sqrt(d > 0.0 ? d : 1.0);
    __syncthreads();
    for (int c = j + tid; c < n; c += BASE_THREADS) S[j * BASE_LD + c] = (c == j) ? sqrt(d > 0.0 ? d : 1.0) : S[j * BASE_LD + c] * rd;
    __syncthreads();
    const int w = n - j - 1;
    for (int idx = tid; idx < w * w; idx += BASE_THREADS) {
      const int r = j + 1 + idx / w, c = j + 1 + idx % w;
      if (r <= c) S[r * BASE_LD + c] -= S[j * BASE_LD + r] * S[j * BASE_LD + c];
    }
    __syncthreads();
  }
  if (write_r)
    for (int idx = tid; idx < n * n; idx += BASE_THREADS) {
      const int r = idx % n, c = idx / n;
      if (r <= c) { if (lower) A[c + (int64_t)r * lda] = S[r * BASE_LD + c]; else A[r + (int64_t)c * lda] = S[r * BASE_LD + c]; }
    }
  // X = R^{-1}: back substitution on all columns at once; X(i,c), c>i lives at S[c][i], X(i,i) in xd.
  // RHS starts as I: lower storage is already zero, the unit diagonal is implicit.
  for (int i = n - 1; i >= 0; --i) {
    const double rinv = 1.0 / S[i * BASE_LD + i];
    __syncthreads();
    if (tid == 0) xd[i] = rinv;                                   // (1 - 0) / R_ii
    for (int c = i + 1 + tid; c < n; c += BASE_THREADS) S[c * BASE_LD + i] *= rinv;
    __syncthreads();
    // rows r < i: B(r, c) -= R(r,i) * X(i,c) for c >= i   (c == i: B(r,i) lives at S[i][r], r<i)
    const int w = n - i;
    for (int idx = tid; idx < i * w; idx += BASE_THREADS) {
      const int r = idx / w, c = i + idx % w;
      const double x = (c == i) ? rinv : S[c * BASE_LD + i];
      S[c * BASE_LD + r] -= S[r * BASE_LD + i] * x;
    }
    __syncthreads();
  }
  for (int idx = tid; idx < n * n; idx += BASE_THREADS) {
    const int r = idx % n, c = idx / n;
    double* out = lower ? &Rinv[c + (int64_t)r * ldri] : &Rinv[r + (int64_t)c * ldri];
    if (r < c) *out = S[c * BASE_LD + r];
    else if (r == c) *out = xd[r];
    else if (zero_lower) *out = 0.0;
  }
  if (tid == 0 && info && s_bad && *info == 0) *info = info_base + s_bad;
}

// ---------------------------------------------------------------- HBM-bound helpers (vectorised where layout allows)
// dst(r,c) = src(r,c) over an m x n window, column-major; mode: 0 all, 1 upper (r<=c) only, 2 lower only,
// 3 upper copy + strict-lower zero, 4 lower copy + strict-upper zero.
__global__ void copy2d_kernel(double* __restrict__ dst, int64_t ldd, const double* __restrict__ src, int64_t lds,
                              int64_t m, int64_t n, int mode) {
  const int64_t c = blockIdx.y;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (int64_t)gridDim.x * blockDim.x) {
    const bool up = r <= c, lo = r >= c;
    if (mode == 0 || (mode == 1 && up) || (mode == 2 && lo)) dst[r + c * ldd] = src[r + c * lds];
    else if (mode == 3) dst[r + c * ldd] = up ? src[r + c * lds] : 0.0;
    else if (mode == 4) dst[r + c * ldd] = lo ? src[r + c * lds] : 0.0;
  }
}

// packed upper (reference uppertri offset X(X+1)/2+Y, src/matrix/structure.h:39) <-> full column-major
__global__ void pack_upper_kernel(double* __restrict__ packed, const double* __restrict__ full, int64_t ld, int n) {
  const int c = blockIdx.y;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r <= c && r < n; r += gridDim.x * blockDim.x)
    packed[(int64_t)c * (c + 1) / 2 + r] = full[r + (int64_t)c * ld];
}
__global__ void unpack_upper_kernel(double* __restrict__ full, int64_t ld, const double* __restrict__ packed, int n, int zero_lower) {
  const int c = blockIdx.y;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    if (r <= c) full[r + (int64_t)c * ld] = packed[(int64_t)c * (c + 1) / 2 + r];
    else if (zero_lower) full[r + (int64_t)c * ld] = 0.0;
  }
}

// y = alpha*x + beta*y, contiguous
__global__ void axpby_kernel(double* __restrict__ y, const double* __restrict__ x, int64_t count, double alpha, double beta) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = alpha * x[i] + (beta == 0.0 ? 0.0 : beta * y[i]);
}

// ---------------------------------------------------------------- drand48 generators (bit-exact with glibc)
constexpr unsigned long long LCG_A = 0x5DEECE66DULL, LCG_C = 0xBULL, LCG_MASK = (1ULL << 48) - 1;

__device__ __forceinline__ void lcg_jump(unsigned long long t, unsigned long long& a_t, unsigned long long& c_t) {
  unsigned long long a = LCG_A, c = LCG_C;
  a_t = 1; c_t = 0;
  while (t) {
    if (t & 1) { c_t = (a * c_t + c) & LCG_MASK; a_t = (a * a_t) & LCG_MASK; }
    c = (a * c + c) & LCG_MASK; a = (a * a) & LCG_MASK;
    t >>= 1;
  }
}

// rect::_distribute_random (structure.hpp:106-129): element t of the rank's sequential stream (column by
// column over the rows x cols block) = drand48 value t after srand48(key).  Each thread jumps ahead once
// and then walks PER_THREAD consecutive elements.
__global__ void fill_random_kernel(double* __restrict__ dst, int64_t ld, int64_t rows, int64_t cols, long long key, int64_t stream_offset) {
  constexpr int PER_THREAD = 16;
  const int64_t total = rows * cols;
  int64_t first = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * PER_THREAD;
  if (first >= total) return;
  unsigned long long x = ((((unsigned long long)key) & 0xFFFFFFFFULL) << 16) | 0x330EULL, a_t, c_t;
  lcg_jump((unsigned long long)(first + stream_offset), a_t, c_t);
  x = (a_t * x + c_t) & LCG_MASK;
  for (int i = 0; i < PER_THREAD && first + i < total; ++i) {
    x = (LCG_A * x + LCG_C) & LCG_MASK;
    const int64_t e = first + i;
    dst[(e % rows) + (e / rows) * ld] = (double)x * (1.0 / 281474976710656.0);
  }
}

// rect::_distribute_symmetric (structure.hpp:69-103): a(gy,gx) = first drand48 after
// srand48(max + globalDimY*min); + diag_add where global row == global col and local row == local col.
__global__ void fill_symmetric_kernel(double* __restrict__ dst, int64_t ld, int64_t rows, int64_t cols, int64_t px, int64_t py,
                                      int64_t grid_x, int64_t grid_y, int64_t global_y, double diag_add) {
  const int64_t c = blockIdx.y;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t gx = px + c * grid_x, gy = py + r * grid_y;
    const long long seed = gx > gy ? gx + global_y * gy : gy + global_y * gx;
    unsigned long long x = ((((unsigned long long)seed) & 0xFFFFFFFFULL) << 16) | 0x330EULL;
    x = (LCG_A * x + LCG_C) & LCG_MASK;
    double v = (double)x * (1.0 / 281474976710656.0);
    if (gx == gy && r == c) v += diag_add;
    dst[r + c * ld] = v;
  }
}

}  // namespace capk
