// Base case of the recursion: R = chol(A) (upper, A = R^T R) and X = R^-1 of an n x n block, n <= 128,
// in ONE CTA.  lapack::engine::_potrf + _trtri as the reference always calls them back to back
// (reference src/lapack/interface.hpp:31-58 <- cacqr.hpp:20-22, cholinv/policy.h:199-201).
//
// Working set in shared memory, as 8 x 8 sub-blocks of 16 x 16:
//   S  : A, overwritten by R, row-major, ld 132;      XB : the 36 upper blocks of X, each TRANSPOSED, ld 20
// (both leading dimensions are 4 mod 16: the (g,t) fragment reads of the FP64 MMA are bank-conflict free
// in either orientation).
//
// What bounds this kernel is one warp's serial pivot chain, so everything is arranged around it (numbers from
// tools/chain_probe.cu and tools/ifetch_probe.cu on a B200: DFMA 8 cycles, MUFU seed 17, shared round trip 32,
// and ~3 cycles per instruction once a warp's straight-line code exceeds ~2000 instructions):
//   S1(b)  warp 0 : finishes A_bb (-= R_{b-1,b}^T R_{b-1,b}) and sweeps the augmented 16 x 32 strip [A_bb | I]:
//                   lanes 0-15 own the columns of A_bb, lanes 16-31 the columns of the identity, so the very
//                   same FMA stream yields R_bb and R_bb^-T.  Per pivot the chain is
//                   store row -> load pivot -> 1/d (seed + 3 FMA) -> update of the next row only;
//                   d^-1/2 is applied to all 16 rows at once after the sweep.
//          warps 1-3, 5-7 : the trailing update left from block row b-1 (all blocks but (b,b)) and
//                   W_i = sum_k X_ik R_kb (i < b);  warp 4 shares warp 0's FP64 pipe and only streams the
//                   finished block row b-1 of R and block column b-1 of X to global memory.
//   S2(b)  all warps, one 16 x 16 item each, in place: R_bj = X_bb^T A_bj (j > b),  X_ib = -W_i X_bb (i < b)
// Two CTA barriers per block column.  The block products run on the FP64 tensor pipe (DMMA.8x8x4).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dmma_gemm.cuh"

namespace capk {

constexpr int BASE_NB = 128;                 // largest block factored by one CTA
constexpr int PB = 16;                       // sub-block
constexpr int PB_NBMAX = BASE_NB / PB;
constexpr int PB_SLD = BASE_NB + 4;          // 4 mod 16
constexpr int PB_XLD = PB + 4;               // 4 mod 16
constexpr int PB_XBLOCKS = PB_NBMAX * (PB_NBMAX + 1) / 2;
constexpr int PB_XSZ = PB * PB_XLD;
constexpr int PB_ROWBUF = PB * 32;           // the 16 un-scaled pivot rows of the sweep
constexpr int BASE_THREADS = 256;
constexpr int BASE_SMEM_BYTES = (BASE_NB * PB_SLD + PB_XBLOCKS * PB_XSZ + PB_ROWBUF + 32) * 8;
static_assert(BASE_SMEM_BYTES <= 232448 - 64, "leaf working set must fit one SM");

#ifdef CAPK_LEAF_PROF
__device__ long long g_leaf_prof[64];
#define LEAF_STAMP(i) do { if (threadIdx.x == 0) g_leaf_prof[i] = clock64(); } while (0)
#else
#define LEAF_STAMP(i) do { } while (0)
#endif

__device__ __forceinline__ int pb_xidx(int bi, int bj) { return bj * (bj + 1) / 2 + bi; }   // bi <= bj

// acc(16x16) += A(16 x 16) * B(16 x 16); acc[i][j][e]: row 8i+g, col 8j+2t+e.  a(r,k), b(k,c) read shared memory.
template <class FA, class FB>
__device__ __forceinline__ void pb_mma16(double (&acc)[2][2][2], FA a, FB b, int g, int t) {
  double af[4][2], bf[4][2];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
#pragma unroll
    for (int i = 0; i < 2; ++i) af[s][i] = a(8 * i + g, 4 * s + t);
#pragma unroll
    for (int j = 0; j < 2; ++j) bf[s][j] = b(4 * s + t, 8 * j + g);
  }
#pragma unroll
  for (int s = 0; s < 4; ++s)
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) dmma884(acc[i][j], af[s][i], bf[s][j]);
}
__device__ __forceinline__ void pb_zero(double (&acc)[2][2][2]) {
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
}
template <class FC>
__device__ __forceinline__ void pb_emit(const double (&acc)[2][2][2], FC c, int g, int t) {
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) c(8 * i + g, 8 * j + 2 * t + e, acc[i][j][e]);
}

// d^-1/2: MUFU.RSQ64H seed and one third-order correction (the library's fast path without its range branch).
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

// Sweep of the augmented strip [A_bb | I] (see the header).  D: block (b,b) of S (upper triangle in, R out);
// Xt: block (b,b) of X, transposed (Xt[k][c] = X(c,k)); U: 16 x 32 row buffer; invs: 32 doubles.
// Lane c < 16 owns column c of A_bb, lane 16 + j column j of the identity; a[r] is row r of that column.
//   step k:  d = U[k][k],  a[r] -= U[k][r] a[k] / d   (r > k),   U[k+1][lane] = a[k+1]
// afterwards, with inv_k = d_k^-1/2:  R(k,c) = a[k] inv_k  (lanes c >= k),   X(j,k) = a[k] inv_k  (lanes 16 + j).
// A pivot that is not a positive normal number counts as non-positive (LAPACK info) and poisons what follows.
// Returns the 1-based row of the first bad pivot inside the block (0 = ok).
__device__ __noinline__ int pb_sweep_potrf(double* D, double* Xt, double* U, double* invs, int lane) {
  double a[PB];
  const int j = lane - PB;
#pragma unroll
  for (int r = 0; r < PB; ++r) a[r] = (lane < PB) ? ((r <= lane) ? D[r * PB_SLD + lane] : 0.0) : ((r == j) ? 1.0 : 0.0);
  U[lane] = a[0];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < PB; ++k) {
    const double* row = U + k * 32;
    const double c = pb_rcp(row[k]);
    if (k + 1 < PB) {
      a[k + 1] = fma(-(row[k + 1] * a[k]), c, a[k + 1]);
      U[(k + 1) * 32 + lane] = a[k + 1];
      __syncwarp();
    }
    const double tk = a[k] * c;
#pragma unroll
    for (int r = k + 2; r < PB; ++r) a[r] = fma(-row[r], tk, a[r]);   // slots below the diagonal are never used
  }
  const double dmine = U[(lane & (PB - 1)) * 32 + (lane & (PB - 1))];
  const bool ok = (dmine >= 2.2250738585072014e-308) && (dmine <= 1.7976931348623157e308);
  const unsigned fail = __ballot_sync(0xffffffffu, !ok) & 0xffffu;
  invs[lane] = pb_rsqrt(dmine);
  __syncwarp();
#pragma unroll
  for (int k = 0; k < PB; ++k) {
    const double v = a[k] * invs[k];
    if (lane < PB) { if (lane >= k) D[k * PB_SLD + lane] = v; }
    else Xt[k * PB_XLD + j] = v;
  }
  return fail ? __ffs(fail) : 0;
}

// Inverse only (R given in D): lane c < 16 solves R^T y = e_c by the same forward sweep.
__device__ __forceinline__ void pb_sweep_inv(const double* D, double* Xt, double* invs, int lane) {
  double y[PB];
  const int c = lane & (PB - 1);
#pragma unroll
  for (int r = 0; r < PB; ++r) y[r] = (r == c) ? 1.0 : 0.0;
  invs[lane] = 1.0 / D[c * PB_SLD + c];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < PB; ++k) {
    y[k] *= invs[k];
#pragma unroll
    for (int r = k + 1; r < PB; ++r) y[r] = fma(-D[k * PB_SLD + r], y[k], y[r]);   // broadcast reads of row k of R
  }
  if (lane < PB) {
#pragma unroll
    for (int k = 0; k < PB; ++k) Xt[k * PB_XLD + c] = y[k];
  }
}

// lane -> (row 0..7, column 0..3) of an 8 x 4 patch such that a half-warp touches 4 rows x 4 columns:
// conflict-free in S (ld 4 mod 16), 64-byte runs in global memory.
__device__ __forceinline__ int patch_row(int lane) { return (lane & 3) | ((lane >> 4) << 2); }
__device__ __forceinline__ int patch_col(int lane) { return (lane >> 2) & 3; }

struct LeafIO {
  double* A; int64_t lda; double* Rinv; int64_t ldri; int n, np, nb, write_r, zero_other, lower;
  __device__ __forceinline__ int64_t oa(int r, int c) const { return lower ? c + (int64_t)r * lda : r + (int64_t)c * lda; }
  __device__ __forceinline__ int64_t ox(int r, int c) const { return lower ? c + (int64_t)r * ldri : r + (int64_t)c * ldri; }
};

// Copy patches (row group rg of 8, column group cg of 4) of the upper triangle of A into S, identity-padded up to np.
// The warp's share is the index range idx = first, first + stride, ... < 16 * 32 with rg = idx >> 5 >= rg_lo and
// rg < rg_hi; up to 8 patches are in flight per lane.
__device__ __forceinline__ void leaf_load(const LeafIO& io, double* S, int first, int stride, int rg_lo, int rg_hi, int lane) {
  const int rl = patch_row(lane), cl = patch_col(lane);
  for (int base = rg_lo * 32 + first; base < rg_hi * 32; base += 8 * stride) {
    double v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int idx = base + u * stride, r = (idx >> 5) * 8 + rl, c = (idx & 31) * 4 + cl;
      v[u] = (r == c) ? 1.0 : 0.0;
      if (idx < rg_hi * 32 && r <= c && c < io.n) v[u] = io.A[io.oa(r, c)];
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int idx = base + u * stride, r = (idx >> 5) * 8 + rl, c = (idx & 31) * 4 + cl;
      if (idx < rg_hi * 32 && r <= c && c < io.np) S[r * PB_SLD + c] = v[u];
    }
  }
}

// Stream the finished block row pb of R and block column pb of X to global memory (zeros below the diagonal
// of that block column when zero_other).  The caller's share: column pairs slot, slot + n_slots, ... of the 8.
__device__ __forceinline__ void leaf_store_panel(const LeafIO& io, const double* S, const double* XB, int pb, int slot, int n_slots, int lane) {
  const int rl = patch_row(lane), cl = patch_col(lane);
  if (io.write_r) {                              // R: rows 16 pb .. 16 pb + 15 = two row groups, columns from 16 pb on
    const int n_cg = (io.np - PB * pb) / 4;
    for (int it = slot; it < 2 * n_cg; it += n_slots) {
      const int r = PB * pb + (it & 1) * 8 + rl, c = PB * pb + (it >> 1) * 4 + cl;
      if (r <= c && c < io.n) io.A[io.oa(r, c)] = S[r * PB_SLD + c];
    }
  }
  const int rr = lane & (PB - 1);
  for (int cp = slot; cp < PB / 2; cp += n_slots) {
    const int cc = 2 * cp + (lane >> 4), c = PB * pb + cc;
    if (c >= io.n) continue;
    for (int i = 0; i <= pb; ++i) {              // X(16 i + rr, c), stored transposed in block (i, pb)
      const int r = PB * i + rr;
      if (r <= c) io.Rinv[io.ox(r, c)] = XB[pb_xidx(i, pb) * PB_XSZ + cc * PB_XLD + rr];
    }
    if (io.zero_other)
      for (int r = PB * pb + rr; r < io.n; r += PB) if (r > c) io.Rinv[io.ox(r, c)] = 0.0;
  }
}

// reads only the upper triangle of A (the lower one, transposed, when `lower`), writes R into that triangle
// when write_r, X = R^-1 into the same triangle of Rinv (other triangle zero-filled when zero_other);
// info (device int, may be null): first non-positive pivot, 1-based, offset by info_base; 0 = ok.
__global__ void __launch_bounds__(BASE_THREADS, 1)
potrf_trtri_base_kernel(double* A, int64_t lda, double* Rinv, int64_t ldri, int n, int* info, int info_base,
                        int write_r, int zero_other, int skip_potrf, int lower) {
  extern __shared__ __align__(16) double sh[];
  double* S = sh;
  double* XB = sh + BASE_NB * PB_SLD;
  double* U = XB + PB_XBLOCKS * PB_XSZ;
  double* invs = U + PB_ROWBUF;
  __shared__ int s_bad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int nb = (n + PB - 1) / PB, np = nb * PB;
  const LeafIO io{A, lda, Rinv, ldri, n, np, nb, write_r, zero_other, lower};
  if (tid == 0) s_bad = 0;
  LEAF_STAMP(0);
  // block row 0 first: the sweep of block 0 starts while the other warps bring in the rest
  leaf_load(io, S, warp, 8, 0, 2, lane);
  __syncthreads();
  LEAF_STAMP(1);

  auto Xblk = [&](int bi, int bj) { return XB + pb_xidx(bi, bj) * PB_XSZ; };     // X(r,c) of the block at [c * PB_XLD + r]
  auto Sblk = [&](int bi, int bj) { return S + (bi * PB) * PB_SLD + bj * PB; };   // A/R(r,c) of the block at [r * PB_SLD + c]

  for (int b = 0; b < nb; ++b) {
    double* Xbb = Xblk(b, b);
    // ---------------------------------------------------------------- S1
    if (warp == 0) {
      double* Dbb = Sblk(b, b);
      if (skip_potrf) {
        pb_sweep_inv(Dbb, Xbb, invs, lane);
      } else {
        if (b > 0) {                                 // last update of A_bb: -= R_{b-1,b}^T R_{b-1,b}
          const double* Rp = Sblk(b - 1, b);
          double acc[2][2][2];
          pb_zero(acc);
          pb_mma16(acc, [&](int r, int k) { return Rp[k * PB_SLD + r]; }, [&](int k, int c) { return Rp[k * PB_SLD + c]; }, g, t);
          pb_emit(acc, [&](int r, int c, double v) { Dbb[r * PB_SLD + c] -= v; }, g, t);
          __syncwarp();
        }
        const int bad = pb_sweep_potrf(Dbb, Xbb, U, invs, lane);
#ifdef CAPK_LEAF_REPEAT
        if (b == 0) {
          LEAF_STAMP(44);
          pb_sweep_potrf(Dbb, Xbb, U, invs, lane);
          LEAF_STAMP(45);
          pb_sweep_potrf(Dbb, Xbb, U, invs, lane);
          LEAF_STAMP(46);
        }
#endif
        if (bad && lane == 0 && s_bad == 0) s_bad = b * PB + bad;
      }
    } else if (b == 0) {
      leaf_load(io, S, warp - 1, 7, 2, np / 8, lane);
    } else if (warp == 4) {
      leaf_store_panel(io, S, XB, b - 1, 0, 1, lane);
    } else {
      const int wk = warp < 4 ? warp - 1 : warp - 2;                    // 6 workers
      const int m = nb - b;                                             // trailing block matrix is m x m
      const int n_trail = skip_potrf ? 0 : m * (m + 1) / 2 - 1;         // every block but (b,b)
      for (int item = wk; item < b + n_trail; item += 6) {
        double acc[2][2][2];
        pb_zero(acc);
        if (item < b) {                              // W_i = sum_{k=i}^{b-1} X_ik R_kb   (longest first)
          const int i = item;
          for (int k = i; k < b; ++k) {
            const double* Xik = Xblk(i, k);
            const double* Rkb = Sblk(k, b);
            pb_mma16(acc, [&](int r, int kk) { return Xik[kk * PB_XLD + r]; }, [&](int kk, int c) { return Rkb[kk * PB_SLD + c]; }, g, t);
          }
          double* dst = Xblk(i, b);
          pb_emit(acc, [&](int r, int c, double v) { dst[c * PB_XLD + r] = v; }, g, t);
        } else {                                     // A_ij -= R_{b-1,i}^T R_{b-1,j}
          int jj = 0, ii = item - b + 1;
          while (ii > jj) { ii -= jj + 1; ++jj; }
          const double* Ri = Sblk(b - 1, b + ii);
          const double* Rj = Sblk(b - 1, b + jj);
          pb_mma16(acc, [&](int r, int k) { return Ri[k * PB_SLD + r]; }, [&](int k, int c) { return Rj[k * PB_SLD + c]; }, g, t);
          double* dst = Sblk(b + ii, b + jj);
          pb_emit(acc, [&](int r, int c, double v) { dst[r * PB_SLD + c] -= v; }, g, t);
        }
      }
    }
    __syncthreads();
    LEAF_STAMP(2 + 2 * b);
    // ---------------------------------------------------------------- S2: one 16 x 16 item per warp, in place
    const int rem = skip_potrf ? 0 : nb - 1 - b;
#ifdef CAPK_LEAF_REPEAT
    for (int rep = 0; rep < 3; ++rep) {
#endif
    for (int item = warp; item < rem + b; item += 8) {
      double acc[2][2][2];
      pb_zero(acc);
      if (item < rem) {                              // R_bj = X_bb^T A_bj
        double* Abj = Sblk(b, b + 1 + item);
        pb_mma16(acc, [&](int r, int k) { return Xbb[r * PB_XLD + k]; }, [&](int k, int c) { return Abj[k * PB_SLD + c]; }, g, t);
        __syncwarp();
        pb_emit(acc, [&](int r, int c, double v) { Abj[r * PB_SLD + c] = v; }, g, t);
      } else {                                       // X_ib = -W_i X_bb
        double* Wi = Xblk(item - rem, b);
        pb_mma16(acc, [&](int r, int k) { return Wi[k * PB_XLD + r]; }, [&](int k, int c) { return Xbb[c * PB_XLD + k]; }, g, t);
        __syncwarp();
        pb_emit(acc, [&](int r, int c, double v) { Wi[c * PB_XLD + r] = -v; }, g, t);
      }
    }
#ifdef CAPK_LEAF_REPEAT
    __syncthreads();
    if (b == 0) LEAF_STAMP(40 + rep);
    }
#endif
    __syncthreads();
    LEAF_STAMP(3 + 2 * b);
  }
  leaf_store_panel(io, S, XB, nb - 1, warp, 8, lane);
  if (tid == 0 && info && s_bad && *info == 0) *info = info_base + s_bad;
#ifdef CAPK_LEAF_PROF
  __syncthreads();
  LEAF_STAMP(30);
#endif
}

}  // namespace capk
