// Base case of the recursion: R = chol(A) (upper, A = R^T R) and X = R^-1 of an n x n block, n <= 128,
// in ONE CTA.  lapack::engine::_potrf + _trtri as the reference always calls them back to back
// (reference src/lapack/interface.hpp:31-58 <- cacqr.hpp:20-22, cholinv/policy.h:199-201).
//
// The block lives in shared memory as 4 x 4 sub-blocks of 32 x 32 (S: A then R, row-major, ld 132;
// XB: the ten upper blocks of X, each stored TRANSPOSED with ld 36 -- both leading dimensions are 4 mod 16
// so that the (g,t) fragment reads of the FP64 MMA are bank-conflict free in either orientation).
//
// The only serial chain is  diag(b) -> R_{b,b+1} -> A_{b+1,b+1} update -> diag(b+1);  everything else
// is scheduled off it (look-ahead).  Per block column b, three CTA barriers:
//   S1  warp 0 : Cholesky of the 32 x 32 diagonal block AND its inverse in the same 32-step sweep
//                (lane c owns column c of A and column c of R^-T; the pivot comes by shuffle, the pivot
//                row by a double-buffered shared row; the inverse reuses the very same row loads)
//       warps 1-7 : the trailing update left over from b-1 (all blocks but (b,b)) and
//                W_i = sum_k X_ik R_kb, i < b   -- neither depends on the diagonal block
//   S2  all warps, half-block items (in place inside one warp): R_bj = X_bb^T A_bj (j > b),  X_ib = -W_i X_bb
//   S3  warps 0-2 : A_{b+1,b+1} -= R_{b,b+1}^T R_{b,b+1}   (the one trailing block the next S1 needs)
// The 16 x 16 x K block products run on the FP64 tensor pipe (DMMA.8x8x4, fragments straight from smem).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dmma_gemm.cuh"

namespace capk {

constexpr int BASE_NB = 128;                 // largest block factored by one CTA
constexpr int PB = 32;                       // sub-block
constexpr int PB_SLD = BASE_NB + 4;          // leading dimension of the 128 x 128 working copy (4 mod 16)
constexpr int PB_XLD = PB + 4;               // leading dimension of one transposed 32 x 32 block of X (4 mod 16)
constexpr int PB_XBLOCKS = 10;               // upper block triangle of a 4 x 4 block matrix
constexpr int PB_XSZ = PB * PB_XLD;
constexpr int BASE_THREADS = 256;
constexpr int BASE_SMEM_BYTES = (BASE_NB * PB_SLD + PB_XBLOCKS * PB_XSZ + 128) * 8;
static_assert(BASE_SMEM_BYTES <= 232448 - 64, "leaf working set must fit one SM");

#ifdef CAPK_LEAF_PROF
__device__ long long g_leaf_prof[64];
#define LEAF_STAMP(i) do { if (threadIdx.x == 0) g_leaf_prof[i] = clock64(); } while (0)
#else
#define LEAF_STAMP(i) do { } while (0)
#endif

__device__ __forceinline__ int pb_xidx(int bi, int bj) { return bj * (bj + 1) / 2 + bi; }   // bi <= bj

// 16x16 quadrant: acc(16x16) += A(16 x [k0,k1)) * B([k0,k1) x 16); acc[i][j][e]: row 8i+g, col 8j+2t+e.
template <class FA, class FB>
__device__ __forceinline__ void pb_mma16(double (&acc)[2][2][2], FA a, FB b, int k0, int k1, int g, int t) {
#pragma unroll 4
  for (int k = k0 + t; k < k1; k += 4) {
    double af[2], bf[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) af[i] = a(8 * i + g, k);
#pragma unroll
    for (int j = 0; j < 2; ++j) bf[j] = b(k, 8 * j + g);
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) dmma884(acc[i][j], af[i], bf[j]);
  }
}
__device__ __forceinline__ void pb_zero(double (&acc)[2][2][2]) {
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
}
// c(r, col, v) for the 16x16 accumulator.
template <class FC>
__device__ __forceinline__ void pb_emit(const double (&acc)[2][2][2], FC c, int g, int t) {
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) c(8 * i + g, 8 * j + 2 * t + e, acc[i][j][e]);
}

// d^-1/2 without the library's out-of-range branch (the caller has already confined d to the normal range):
// MUFU.RSQ64H seed and one third-order correction, the sequence the library's fast path uses.
__device__ __forceinline__ double pb_rsqrt(double d) {
  double y0;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(d));
  const double e = fma(-(y0 * y0), d, 1.0);
  return fma(fma(e, 0.375, 0.5), y0 * e, y0);
}

// 1/d from the MUFU.RCP64H seed (2^-20.5) by one third-order step: 3 dependent FMAs, relative error ~2^-52.
__device__ __forceinline__ double pb_rcp(double d) {
  double c0;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(c0) : "d"(d));
  const double e = fma(-d, c0, 1.0);
  return fma(c0, fma(e, e, e), c0);
}

// Warp-level Cholesky of the 32 x 32 diagonal block at D (ld PB_SLD, upper triangle, R written back in
// place) fused with its inverse: Xt (ld PB_XLD) receives X = R^-1 transposed, i.e. Xt[k][c] = X(c,k).
//
// Two halves of 16 pivots (a real loop: the body must stay inside the instruction cache -- one warp running
// > 2000 straight-line instructions is fetch-bound at ~3 cycles each, measured by tools/ifetch_probe.cu).
// In a half, lane c holds rows l0..l0+15 of column c: a[r] = A(l0+r, c), so that the sweep factors the
// 16 x 16 diagonal block in lanes l0..l0+15 and, in the first half, solves the 16 x 16 panel R12 in lanes
// 16..31 on the way.  y[r] = (R^-T)(l0+r, c) accumulates the inverse of the diagonal block from the same
// pivot-row loads.  Step k, pivot d, un-scaled pivot row row[]:
//     a[r] -= row[r] a[k] / d,   y[r] -= row[r] y[k] / d      (r > k)
// The serial chain is shuffle(d) -> 1/d (seed + 3 FMA) -> the update of slot k+1 alone -> next shuffle;
// d^-1/2 (only needed to scale the finished row) and every other slot ride underneath it.
// Between the halves A22 -= R12^T R12, and at the end X12 = -X11 R12 X22, on the FP64 tensor pipe.
// A pivot that is not a positive normal number counts as non-positive (LAPACK info) and poisons what follows.
// buf: 128 doubles of warp-private shared memory.  Returns the 1-based index of the first bad pivot (0 = ok).
__device__ __forceinline__ int pb_diag_potrf_inv(double* D, double* Xt, double* buf, int lane, int g, int t) {
  constexpr int H = PB / 2;
  int bad = 0;
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    const int l0 = H * half;
    double* Dh = D + l0 * PB_SLD;
    double a[H], y[H];
    LEAF_STAMP(20 + 5 * half);
#pragma unroll
    for (int r = 0; r < H; ++r) {
      a[r] = (l0 + r <= lane) ? Dh[r * PB_SLD + lane] : 0.0;
      y[r] = (l0 + r == lane) ? 1.0 : 0.0;
    }
    buf[lane] = a[0];
    double d = __shfl_sync(0xffffffffu, a[0], l0);
    __syncwarp();
    LEAF_STAMP(21 + 5 * half);
    double dmine = 1.0;                             // the pivot of this lane's own column
#pragma unroll
    for (int k = 0; k < H; ++k) {
      const double* row = buf + (k % 3) * PB + l0;        // row[r]: pivot-row entry of column l0 + r
      const double c = pb_rcp(d);
      double dn = 0.0;
      if (k + 1 < H) {
        a[k + 1] = fma(-(row[k + 1] * a[k]), c, a[k + 1]);
        buf[((k + 1) % 3) * PB + lane] = a[k + 1];
        dn = __shfl_sync(0xffffffffu, a[k + 1], l0 + k + 1);
        __syncwarp();
      }
      dmine = (lane == l0 + k) ? d : dmine;
      const double tk = a[k] * c, uk = y[k] * c;
      if (k + 1 < H) y[k + 1] = fma(-row[k + 1], uk, y[k + 1]);
#pragma unroll
      for (int r = k + 2; r < H; ++r) {
        const double rr = row[r];
        a[r] = fma(-rr, tk, a[r]);                // slots below the diagonal of the block are never used
        y[r] = fma(-rr, uk, y[r]);                // zero while lane > l0 + k
      }
      d = dn;
    }
    // scaling by d^-1/2, all 16 pivots at once (lane l0 + k owns pivot k): R(l0+k, c) = a[k] inv_k, Y(k, c) = y[k] inv_k
    {
      const bool mine = lane >= l0 && lane < l0 + H;
      const bool ok = !mine || ((dmine >= 2.2250738585072014e-308) && (dmine <= 1.7976931348623157e308));
      const unsigned fail = __ballot_sync(0xffffffffu, !ok);
      if (fail && bad == 0) bad = __ffs(fail);            // lane index + 1 = 1-based row inside the block
      double* invs = buf + 3 * PB;
      invs[lane] = pb_rsqrt(dmine);
      __syncwarp();
#pragma unroll
      for (int k = 0; k < H; ++k) {
        const double inv = invs[l0 + k];
        if (lane >= l0 + k) Dh[k * PB_SLD + lane] = a[k] * inv;
        y[k] *= inv;
      }
    }
    LEAF_STAMP(22 + 5 * half);
    if (half == 0 || lane >= H) {                 // first half also writes the zeros of X below the diagonal
#pragma unroll
      for (int k = 0; k < H; ++k) Xt[(l0 + k) * PB_XLD + lane] = y[k];
    }
    __syncwarp();
    LEAF_STAMP(23 + 5 * half);
    if (half == 0) {                              // A22 -= R12^T R12
      double acc[2][2][2];
      pb_zero(acc);
      const double* R12 = D + H;
      pb_mma16(acc, [&](int r, int k) { return R12[k * PB_SLD + r]; }, [&](int k, int cc) { return R12[k * PB_SLD + cc]; }, 0, H, g, t);
      double* A22 = D + H * PB_SLD + H;
      pb_emit(acc, [&](int r, int cc, double v) { A22[r * PB_SLD + cc] -= v; }, g, t);
      __syncwarp();
    }
    LEAF_STAMP(24 + 5 * half);
  }
  {                                               // X12 = -X11 (R12 X22), in the X12 slot Xt[(16 + c)][r]
    const double* R12 = D + H;
    double* X12 = Xt + H * PB_XLD;
    double acc[2][2][2];
    pb_zero(acc);
    pb_mma16(acc, [&](int r, int k) { return R12[r * PB_SLD + k]; }, [&](int k, int cc) { return X12[cc * PB_XLD + H + k]; }, 0, H, g, t);
    pb_emit(acc, [&](int r, int cc, double v) { X12[cc * PB_XLD + r] = v; }, g, t);
    __syncwarp();
    pb_zero(acc);
    pb_mma16(acc, [&](int r, int k) { return Xt[k * PB_XLD + r]; }, [&](int k, int cc) { return X12[cc * PB_XLD + k]; }, 0, H, g, t);
    __syncwarp();
    pb_emit(acc, [&](int r, int cc, double v) { X12[cc * PB_XLD + r] = -v; }, g, t);
  }
  LEAF_STAMP(30);
  return bad;
}

// Inverse only (R given in D): the same forward sweep without the factorisation.
__device__ __forceinline__ void pb_diag_inv(const double* D, double* Xt, double* buf, int lane) {
  double y[PB];
#pragma unroll
  for (int r = 0; r < PB; ++r) y[r] = (r == lane) ? 1.0 : 0.0;
  buf[lane] = 1.0 / D[lane * PB_SLD + lane];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < PB; ++k) {
    y[k] *= buf[k];
#pragma unroll
    for (int r = k + 1; r < PB; ++r) y[r] = fma(-D[k * PB_SLD + r], y[k], y[r]);   // broadcast reads of row k of R
  }
#pragma unroll
  for (int k = 0; k < PB; ++k) Xt[k * PB_XLD + lane] = y[k];
}

// Upper-triangle walk shared by the load and the store: warp w visits the 8-row x 4-column patches of
// column-group pairs {w, 15-w} (17 row groups x 2 column groups = 34 patches per warp, balanced), lane ->
// (row 0..7, column 0..3) inside a patch so that a half-warp touches 4 rows x 4 columns: conflict-free in S
// (ld 4 mod 16) and in the transposed X blocks, 64-byte runs in global memory.
struct PatchWalk {
  int rl, cl, w;
  __device__ __forceinline__ PatchWalk(int warp, int lane) : rl((lane & 3) | ((lane >> 4) << 2)), cl((lane >> 2) & 3), w(warp) {}
  // patch p in [0,34): element (r, c) of this lane
  __device__ __forceinline__ void at(int p, int& r, int& c) const {
    const int q = p >> 1;
    const int cg2 = (q <= w) ? w : 15 - w, rg = (q <= w) ? q : q - w - 1;
    r = rg * 8 + rl;
    c = (cg2 * 2 + (p & 1)) * 4 + cl;
  }
};

// reads only the upper triangle of A (the lower one, transposed, when `lower`), writes R into that triangle
// when write_r, X = R^-1 into the same triangle of Rinv (other triangle zero-filled when zero_other);
// info (device int, may be null): first non-positive pivot, 1-based, offset by info_base; 0 = ok.
__global__ void __launch_bounds__(BASE_THREADS, 1)
potrf_trtri_base_kernel(double* A, int64_t lda, double* Rinv, int64_t ldri, int n, int* info, int info_base,
                        int write_r, int zero_other, int skip_potrf, int lower,
                        int leaves_per_xblock = 1 << 30, int64_t xblock_stride = 0) {
  // batched form (independent diagonal blocks, e.g. every leaf of a triangular inverse in one launch): CTA b takes
  // the BASE_NB block at (b, b) * BASE_NB of an n x n matrix; its inverse goes to the same place of Rinv, or, when
  // Rinv is a row of compact square blocks of leaves_per_xblock leaves each, to block b / leaves_per_xblock.
  if (blockIdx.x) {
    const int b = blockIdx.x;
    A += (int64_t)b * BASE_NB * (lda + 1);
    Rinv += (int64_t)(b / leaves_per_xblock) * xblock_stride + (int64_t)(b % leaves_per_xblock) * BASE_NB * (ldri + 1);
    info_base += b * BASE_NB;
    n -= b * BASE_NB;
  }
  n = n < BASE_NB ? n : BASE_NB;
  extern __shared__ __align__(16) double sh[];
  double* S = sh;
  double* XB = sh + BASE_NB * PB_SLD;
  double* buf = XB + PB_XBLOCKS * PB_XSZ;
  __shared__ int s_bad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int nb = (n + PB - 1) / PB, np = nb * PB;
  const PatchWalk walk(warp, lane);
  if (tid == 0) s_bad = 0;
  LEAF_STAMP(0);
  // ---- working copy of the upper triangle, padded with the identity up to a multiple of 32; all loads in flight
  {
    double v[34];
#pragma unroll
    for (int p = 0; p < 34; ++p) {
      int r, c; walk.at(p, r, c);
      v[p] = (r == c) ? 1.0 : 0.0;
      if (r <= c && c < n) v[p] = lower ? A[c + (int64_t)r * lda] : A[r + (int64_t)c * lda];
    }
#pragma unroll
    for (int p = 0; p < 34; ++p) {
      int r, c; walk.at(p, r, c);
      if (c < np && r <= c) S[r * PB_SLD + c] = v[p];
    }
  }
  __syncthreads();
  LEAF_STAMP(1);

  auto Xblk = [&](int bi, int bj) { return XB + pb_xidx(bi, bj) * PB_XSZ; };     // X(r,c) of the block at [c * PB_XLD + r]
  auto Sblk = [&](int bi, int bj) { return S + (bi * PB) * PB_SLD + bj * PB; };   // A/R(r,c) of the block at [r * PB_SLD + c]

  for (int b = 0; b < nb; ++b) {
    double* Xbb = Xblk(b, b);
    // ---- S1: diagonal block on warp 0; left-over trailing update of b-1 and W_i(b) on the others
    if (warp == 0) {
      if (skip_potrf) {
        pb_diag_inv(Sblk(b, b), Xbb, buf, lane);
      } else {
        const int bad = pb_diag_potrf_inv(Sblk(b, b), Xbb, buf, lane, g, t);
        if (bad && lane == 0 && s_bad == 0) s_bad = b * PB + bad;
      }
    } else if (b > 0) {
      const int m = nb - b;                                            // trailing block matrix is m x m
      const int n_trail = skip_potrf ? 0 : (m * (m + 1) / 2 - 1) * 4;  // quadrants of all blocks but (b,b)
      const int n_items = n_trail + b * 4;
      // warp 4 shares warp 0's sub-partition (and its FP64 pipe): it sits S1 out
      for (int item = (warp < 4 ? warp - 1 : warp - 2); warp != 4 && item < n_items; item += 6) {
        const int q = item & 3, r0 = (q >> 1) * 16, c0 = (q & 1) * 16;
        double acc[2][2][2];
        pb_zero(acc);
        if (item < n_trail) {                      // A_ij -= R_{b-1,i}^T R_{b-1,j}
          int jj = 0, ii = (item >> 2) + 1;
          while (ii > jj) { ii -= jj + 1; ++jj; }
          if (ii == jj && q == 2) continue;        // strictly lower quadrant of a diagonal block
          const double* Ri = Sblk(b - 1, b + ii) + r0;
          const double* Rj = Sblk(b - 1, b + jj) + c0;
          pb_mma16(acc, [&](int r, int k) { return Ri[k * PB_SLD + r]; }, [&](int k, int c) { return Rj[k * PB_SLD + c]; }, 0, PB, g, t);
          double* dst = Sblk(b + ii, b + jj) + r0 * PB_SLD + c0;
          pb_emit(acc, [&](int r, int c, double v) { dst[r * PB_SLD + c] -= v; }, g, t);
        } else {                                   // W_i = sum_{k=i}^{b-1} X_ik R_kb
          const int i = (item - n_trail) >> 2;
          for (int k = i; k < b; ++k) {
            const double* Xik = Xblk(i, k) + r0;
            const double* Rkb = Sblk(k, b) + c0;
            pb_mma16(acc, [&](int r, int kk) { return Xik[kk * PB_XLD + r]; }, [&](int kk, int c) { return Rkb[kk * PB_SLD + c]; },
                     k == i ? r0 : 0, PB, g, t);   // X_ii is upper triangular
          }
          double* dst = Xblk(i, b) + c0 * PB_XLD + r0;
          pb_emit(acc, [&](int r, int c, double v) { dst[c * PB_XLD + r] = v; }, g, t);
        }
      }
    }
    __syncthreads();
    LEAF_STAMP(2 + 3 * b);
    // ---- S2: half-block items, each in place inside one warp (all fragment loads precede the stores)
    const int rem = skip_potrf ? 0 : nb - 1 - b;
    for (int item = warp; item < 2 * (rem + b); item += 8) {
      const int half = item & 1;
      double acc[2][2][2][2];
      pb_zero(acc[0]); pb_zero(acc[1]);
      if (item < 2 * rem) {                        // R_bj = X_bb^T A_bj, columns [16 half, 16 half + 16)
        double* Abj = Sblk(b, b + 1 + (item >> 1)) + 16 * half;
        auto xa0 = [&](int r, int k) { return Xbb[r * PB_XLD + k]; };            // X_bb(k, r), rows 0..15: k < 16 suffices
        auto xa1 = [&](int r, int k) { return Xbb[(16 + r) * PB_XLD + k]; };
        auto bb = [&](int k, int c) { return Abj[k * PB_SLD + c]; };
        pb_mma16(acc[0], xa0, bb, 0, 16, g, t);
        pb_mma16(acc[1], xa1, bb, 0, PB, g, t);
        __syncwarp();
        pb_emit(acc[0], [&](int r, int c, double v) { Abj[r * PB_SLD + c] = v; }, g, t);
        pb_emit(acc[1], [&](int r, int c, double v) { Abj[(16 + r) * PB_SLD + c] = v; }, g, t);
      } else {                                     // X_ib = -W_i X_bb, rows [16 half, 16 half + 16)
        double* Wi = Xblk((item - 2 * rem) >> 1, b) + 16 * half;
        auto wa = [&](int r, int k) { return Wi[k * PB_XLD + r]; };
        auto xb0 = [&](int k, int c) { return Xbb[c * PB_XLD + k]; };            // X_bb(k, c), columns 0..15: k < 16
        auto xb1 = [&](int k, int c) { return Xbb[(16 + c) * PB_XLD + k]; };
        pb_mma16(acc[0], wa, xb0, 0, 16, g, t);
        pb_mma16(acc[1], wa, xb1, 0, PB, g, t);
        __syncwarp();
        pb_emit(acc[0], [&](int r, int c, double v) { Wi[c * PB_XLD + r] = -v; }, g, t);
        pb_emit(acc[1], [&](int r, int c, double v) { Wi[(16 + c) * PB_XLD + r] = -v; }, g, t);
      }
    }
    __syncthreads();
    LEAF_STAMP(3 + 3 * b);
    // ---- S3: the one trailing block the next diagonal step needs
    if (rem > 0) {
      if (warp < 3) {
        const int q = warp == 2 ? 3 : warp, r0 = (q >> 1) * 16, c0 = (q & 1) * 16;
        double acc[2][2][2];
        pb_zero(acc);
        const double* Ri = Sblk(b, b + 1) + r0;
        const double* Rj = Sblk(b, b + 1) + c0;
        pb_mma16(acc, [&](int r, int k) { return Ri[k * PB_SLD + r]; }, [&](int k, int c) { return Rj[k * PB_SLD + c]; }, 0, PB, g, t);
        double* dst = Sblk(b + 1, b + 1) + r0 * PB_SLD + c0;
        pb_emit(acc, [&](int r, int c, double v) { dst[r * PB_SLD + c] -= v; }, g, t);
      }
      __syncthreads();
    }
    LEAF_STAMP(4 + 3 * b);
  }

  // ---- results: upper triangle by patches, the other triangle zero-filled column by column
#pragma unroll 2
  for (int p = 0; p < 34; ++p) {
    int r, c; walk.at(p, r, c);
    if (c < n && r <= c) {
      const int64_t oa = lower ? c + (int64_t)r * lda : r + (int64_t)c * lda;
      const int64_t ox = lower ? c + (int64_t)r * ldri : r + (int64_t)c * ldri;
      if (write_r) A[oa] = S[r * PB_SLD + c];
      Rinv[ox] = Xblk(r / PB, c / PB)[(c % PB) * PB_XLD + (r % PB)];
    }
  }
  if (zero_other) {
    for (int c = warp; c < n; c += 8)
      for (int r = c + 1 + lane; r < n; r += 32) Rinv[lower ? c + (int64_t)r * ldri : r + (int64_t)c * ldri] = 0.0;
  }
  if (tid == 0 && info && s_bad && *info == 0) *info = info_base + s_bad;
#ifdef CAPK_LEAF_PROF
  __syncthreads();
  LEAF_STAMP(14);
#endif
}

}  // namespace capk
