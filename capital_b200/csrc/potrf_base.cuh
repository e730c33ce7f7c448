// Base case of the recursion: R = chol(A) (upper, A = R^T R) and X = R^-1 of an n x n block, n <= 128,
// in ONE CTA.  lapack::engine::_potrf + _trtri as the reference always calls them back to back
// (reference src/lapack/interface.hpp:31-58 <- cacqr.hpp:20-22, cholinv/policy.h:199-201).
//
// The block lives in shared memory as 4 x 4 sub-blocks of 32 x 32.  Per block column b:
//   phase 1  warp 0: register-blocked Cholesky of the 32 x 32 diagonal block (lane c owns column c; the
//            pivot row is broadcast through a 32-double shared buffer, one __syncwarp per column) and
//            its inverse by column-oriented back substitution (lane c solves R x = e_c);
//   phase 2  one warp per block: row panel R_bj = X_bb^T A_bj (j > b)   ||   W_i = sum_k X_ik R_kb (i < b)
//   phase 3  one warp per block: trailing A_ij -= R_bi^T R_bj (b < i <= j)  ||  X_ib = -W_i X_bb
// The 32 x 32 x 32 block products run on the FP64 tensor pipe (DMMA.8x8x4, fragments straight from
// shared memory).  12 CTA barriers in total instead of ~6 per matrix column.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dmma_gemm.cuh"

namespace capk {

constexpr int BASE_NB = 128;                 // largest block factored by one CTA
constexpr int PB = 32;                       // sub-block
constexpr int PB_SLD = BASE_NB + 1;          // leading dimension of the 128 x 128 working copy
constexpr int PB_XLD = PB + 1;               // leading dimension of one 32 x 32 block of X
constexpr int PB_XBLOCKS = 10;               // upper block triangle of a 4 x 4 block matrix
constexpr int BASE_THREADS = 256;           // 8 warps: never more than 6 independent block products per phase
constexpr int BASE_SMEM_BYTES = (BASE_NB * PB_SLD + PB_XBLOCKS * PB * PB_XLD + 64) * 8;

__device__ __forceinline__ int pb_xidx(int bi, int bj) { return bj * (bj + 1) / 2 + bi; }   // bi <= bj

// acc(32x32) += A(32x32) * B(32x32); a(r,k), b(k,c) read shared memory.  acc[i][j][e]: row 8i+g, col 8j+2t+e.
template <class FA, class FB>
__device__ __forceinline__ void pb_mma32(double (&acc)[4][4][2], FA a, FB b, int g, int t) {
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    double af[4], bf[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) af[i] = a(8 * i + g, 4 * s + t);
#pragma unroll
    for (int j = 0; j < 4; ++j) bf[j] = b(4 * s + t, 8 * j + g);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) dmma884(acc[i][j], af[i], bf[j]);
  }
}
// 16x16 quadrant: acc(16x16) += A(16xK) * B(Kx16); acc[i][j][e]: row 8i+g, col 8j+2t+e.
template <class FA, class FB>
__device__ __forceinline__ void pb_mma16(double (&acc)[2][2][2], FA a, FB b, int K, int g, int t) {
#pragma unroll 8
  for (int s = 0; s < K / 4; ++s) {
    double af[2], bf[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) af[i] = a(8 * i + g, 4 * s + t);
#pragma unroll
    for (int j = 0; j < 2; ++j) bf[j] = b(4 * s + t, 8 * j + g);
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) dmma884(acc[i][j], af[i], bf[j]);
  }
}
__device__ __forceinline__ void pb_zero(double (&acc)[4][4][2]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
}
template <class FC>
__device__ __forceinline__ void pb_store(const double (&acc)[4][4][2], FC c, int g, int t) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) c(8 * i + g, 8 * j + 2 * t + e, acc[i][j][e]);
}

// Warp-level Cholesky of the 32 x 32 diagonal block at D (leading dimension PB_SLD, upper triangle) and its
// inverse into X (dense 32 x 32, leading dimension PB_XLD, zeros below the diagonal).
// buf: 64 doubles of warp-private shared memory.  Returns the 1-based index of the first bad pivot (0 = ok).
__device__ __forceinline__ int pb_diag_block(double* D, double* X, double* buf, int lane, bool do_potrf) {
  double a[PB];
  int bad = 0;
#pragma unroll
  for (int r = 0; r < PB; ++r) a[r] = (r <= lane) ? D[r * PB_SLD + lane] : 0.0;   // lane = column
  if (do_potrf) {
#pragma unroll
    for (int k = 0; k < PB; ++k) {
      double* row = buf + (k & 1) * PB;           // double-buffered pivot row: one __syncwarp per column
      row[lane] = a[k];
      __syncwarp();
      double d = row[k];
      if (!(d > 0.0)) { if (bad == 0) bad = k + 1; d = 1.0; }
      const double inv = rsqrt(d);
      const double rk = a[k] * inv;               // R(k, lane) for lane > k
      if (lane == k) a[k] = d * inv; else a[k] = rk;
      const double tk = rk * inv;
#pragma unroll
      for (int r = k + 1; r < PB; ++r) a[r] = fma(-row[r], tk, a[r]);   // A(r,c) -= R(k,r) R(k,c); slots r > lane are never used
    }
#pragma unroll
    for (int r = 0; r < PB; ++r) if (r <= lane) D[r * PB_SLD + lane] = a[r];
  }
  // inverse: lane c solves R x = e_c, column-oriented back substitution (independent FMAs per step)
  double diag = 1.0;
#pragma unroll
  for (int r = 0; r < PB; ++r) if (r == lane) diag = a[r];   // R(lane, lane) without dynamic register indexing
  buf[lane] = 1.0 / diag;
  __syncwarp();
  double x[PB];
#pragma unroll
  for (int r = 0; r < PB; ++r) x[r] = (r == lane) ? 1.0 : 0.0;
#pragma unroll
  for (int k = PB - 1; k >= 0; --k) {
    x[k] *= buf[k];
#pragma unroll
    for (int r = 0; r < k; ++r) x[r] = fma(-D[r * PB_SLD + k], x[k], x[r]);       // broadcast reads of column k of R
  }
#pragma unroll
  for (int r = 0; r < PB; ++r) X[r * PB_XLD + lane] = x[r];
  return bad;
}

// Interface identical to the column-by-column kernel it replaces:
// reads only the upper triangle of A (the lower one, transposed, when `lower`), writes R into that triangle
// when write_r, X = R^-1 into the same triangle of Rinv (other triangle zero-filled when zero_other);
// info (device int, may be null): first non-positive pivot, 1-based, offset by info_base; 0 = ok.
__global__ void __launch_bounds__(BASE_THREADS, 1)
potrf_trtri_base_kernel(double* A, int64_t lda, double* Rinv, int64_t ldri, int n, int* info, int info_base,
                        int write_r, int zero_other, int skip_potrf, int lower) {
  extern __shared__ __align__(16) double sh[];
  double* S = sh;
  double* XB = sh + BASE_NB * PB_SLD;
  double* buf = XB + PB_XBLOCKS * PB * PB_XLD;
  __shared__ int s_bad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int nb = (n + PB - 1) / PB, np = nb * PB;
  if (tid == 0) s_bad = 0;
  // working copy, padded with the identity up to a multiple of 32
  for (int idx = tid; idx < np * np; idx += BASE_THREADS) {
    const int r = idx % np, c = idx / np;
    double v = 0.0;
    if (r <= c) {
      if (c < n) v = lower ? A[c + (int64_t)r * lda] : A[r + (int64_t)c * lda];
      else v = (r == c) ? 1.0 : 0.0;
    }
    S[r * PB_SLD + c] = v;
  }
  __syncthreads();

  for (int b = 0; b < nb; ++b) {
    double* Xbb = XB + pb_xidx(b, b) * PB * PB_XLD;
    // ---- phase 1: diagonal block
    if (warp == 0) {
      const int bad = pb_diag_block(S + (b * PB) * PB_SLD + b * PB, Xbb, buf, lane, !skip_potrf);
      if (bad && lane == 0 && s_bad == 0) s_bad = b * PB + bad;
    }
    __syncthreads();
    // ---- phases 2 and 3: 32x32 block products, each split into four 16x16 quadrants = one warp-item each.
    // Results stay in registers until every warp has finished reading (several products are in place).
    const int rem = nb - 1 - b;
    for (int phase = 2; phase <= 3; ++phase) {
      const int n_main = skip_potrf ? 0 : (phase == 2 ? rem : rem * (rem + 1) / 2);   // panel blocks / trailing blocks
      const int n_items = (n_main + b) * 4;
      double acc[3][2][2][2];
#pragma unroll
      for (int round = 0; round < 3; ++round) {
        const int item = warp + 8 * round;
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int j = 0; j < 2; ++j) acc[round][i][j][0] = acc[round][i][j][1] = 0.0;
        if (item >= n_items) continue;
        const int blk = item >> 2, r0 = ((item >> 1) & 1) * 16, c0 = (item & 1) * 16;
        if (blk < n_main && phase == 2) {          // R_bj = X_bb^T A_bj
          const double* Abj = S + (b * PB) * PB_SLD + (b + 1 + blk) * PB;
          pb_mma16(acc[round], [&](int r, int k) { return Xbb[k * PB_XLD + r0 + r]; }, [&](int k, int c) { return Abj[k * PB_SLD + c0 + c]; }, PB, g, t);
        } else if (blk < n_main) {                 // A_ij -= R_bi^T R_bj
          int jj = 0, ii = blk;
          while (ii > jj) { ii -= jj + 1; ++jj; }
          const double* Rbi = S + (b * PB) * PB_SLD + (b + 1 + ii) * PB;
          const double* Rbj = S + (b * PB) * PB_SLD + (b + 1 + jj) * PB;
          pb_mma16(acc[round], [&](int r, int k) { return Rbi[k * PB_SLD + r0 + r]; }, [&](int k, int c) { return Rbj[k * PB_SLD + c0 + c]; }, PB, g, t);
        } else if (phase == 2) {                   // W_i = sum_k X_ik R_kb
          const int i = blk - n_main;
          for (int k = i; k < b; ++k) {
            const double* Xik = XB + pb_xidx(i, k) * PB * PB_XLD;
            const double* Rkb = S + (k * PB) * PB_SLD + b * PB;
            pb_mma16(acc[round], [&](int r, int kk) { return Xik[(r0 + r) * PB_XLD + kk]; }, [&](int kk, int c) { return Rkb[kk * PB_SLD + c0 + c]; }, PB, g, t);
          }
        } else {                                   // X_ib = -W_i X_bb
          const double* Wi = XB + pb_xidx(blk - n_main, b) * PB * PB_XLD;
          pb_mma16(acc[round], [&](int r, int k) { return Wi[(r0 + r) * PB_XLD + k]; }, [&](int k, int c) { return Xbb[k * PB_XLD + c0 + c]; }, PB, g, t);
        }
      }
      __syncthreads();
#pragma unroll
      for (int round = 0; round < 3; ++round) {
        const int item = warp + 8 * round;
        if (item >= n_items) continue;
        const int blk = item >> 2, r0 = ((item >> 1) & 1) * 16, c0 = (item & 1) * 16;
        double* dst; int ld; double sign = 1.0; bool accumulate = false;
        if (blk < n_main && phase == 2) { dst = S + (b * PB) * PB_SLD + (b + 1 + blk) * PB; ld = PB_SLD; }
        else if (blk < n_main) {
          int jj = 0, ii = blk;
          while (ii > jj) { ii -= jj + 1; ++jj; }
          dst = S + ((b + 1 + ii) * PB) * PB_SLD + (b + 1 + jj) * PB; ld = PB_SLD; sign = -1.0; accumulate = true;
        } else { dst = XB + pb_xidx(blk - n_main, b) * PB * PB_XLD; ld = PB_XLD; sign = (phase == 2) ? 1.0 : -1.0; }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              double* q = dst + (r0 + 8 * i + g) * ld + c0 + 8 * j + 2 * t + e;
              *q = accumulate ? *q + sign * acc[round][i][j][e] : sign * acc[round][i][j][e];
            }
      }
      __syncthreads();
    }
  }

  // ---- results
  for (int idx = tid; idx < n * n; idx += BASE_THREADS) {
    const int r = idx % n, c = idx / n;
    double* xo = lower ? &Rinv[c + (int64_t)r * ldri] : &Rinv[r + (int64_t)c * ldri];
    if (r <= c) {
      if (write_r) { if (lower) A[c + (int64_t)r * lda] = S[r * PB_SLD + c]; else A[r + (int64_t)c * lda] = S[r * PB_SLD + c]; }
      *xo = XB[pb_xidx(r / PB, c / PB) * PB * PB_XLD + (r % PB) * PB_XLD + (c % PB)];
    } else if (zero_other) {
      *xo = 0.0;
    }
  }
  if (tid == 0 && info && s_bad && *info == 0) *info = info_base + s_bad;
}

}  // namespace capk
