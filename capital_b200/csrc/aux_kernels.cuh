// HBM-bound helper kernels of the path (pack/unpack, copies, masks, axpby, generators); the small-block
// Cholesky + triangular inverse lives in potrf_leaf.cuh.
//   pack/unpack/copy          serialize<S1,S2>::invoke        (src/matrix/serialize.hpp:12-150)
//   mask                      util::remove_triangle           (src/util/util.hpp:267-293)
//   axpby                     summa.hpp:33,153
//   generators                rect::_distribute_random/_symmetric (src/matrix/structure.hpp:69-129)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>

#include "potrf_leaf.cuh"

namespace capk {

// ---------------------------------------------------------------- HBM-bound helpers
// dst(r,c) = src(r,c) over an m x n window, column-major; MODE: 0 all, 1 upper (r<=c) only, 2 lower only,
// 3 upper copy + strict-lower zero, 4 lower copy + strict-upper zero.  VEC: both bases 16-byte aligned and both leading
// dimensions even -- rows travel in pairs (one 16-byte access each way per thread and iteration).  Columns are walked with
// a stride of gridDim.y, so any column count fits the 65535 limit of that grid dimension.
constexpr int COPY_COLS = 8;      // columns per thread and pass: eight independent 16-byte loads in flight before the first store
template <int MODE, bool VEC>
__global__ void copy2d_kernel(double* __restrict__ dst, int64_t ldd, const double* __restrict__ src, int64_t lds, int64_t m, int64_t n) {
  auto keep = [](int64_t r, int64_t c) { return MODE == 0 || ((MODE == 1 || MODE == 3) ? r <= c : r >= c); };
  for (int64_t c0 = (int64_t)blockIdx.y * COPY_COLS; c0 < n; c0 += (int64_t)gridDim.y * COPY_COLS) {
    if (VEC) {
      const int64_t pairs = m >> 1;
      for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < pairs; q += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = 2 * q;
        double2 v[COPY_COLS];
#pragma unroll
        for (int j = 0; j < COPY_COLS; ++j) {
          const int64_t c = c0 + j;
          v[j] = make_double2(0.0, 0.0);
          if (c < n && (keep(r, c) || keep(r + 1, c))) v[j] = *reinterpret_cast<const double2*>(src + c * lds + r);
        }
#pragma unroll
        for (int j = 0; j < COPY_COLS; ++j) {
          const int64_t c = c0 + j;
          if (c >= n) continue;
          double* d = dst + c * ldd + r;
          const bool k0 = keep(r, c), k1 = keep(r + 1, c);
          if (MODE == 1 || MODE == 2) {                   // masked entries of dst stay as they are
            if (k0 && k1) *reinterpret_cast<double2*>(d) = v[j];
            else { if (k0) d[0] = v[j].x; if (k1) d[1] = v[j].y; }
          } else {
            if (!k0) v[j].x = 0.0;
            if (!k1) v[j].y = 0.0;
            *reinterpret_cast<double2*>(d) = v[j];
          }
        }
      }
      if ((m & 1) && blockIdx.x == 0 && threadIdx.x < COPY_COLS && c0 + threadIdx.x < n) {
        const int64_t r = m - 1, c = c0 + threadIdx.x;
        if (keep(r, c)) dst[c * ldd + r] = src[c * lds + r]; else if (MODE >= 3) dst[c * ldd + r] = 0.0;
      }
    } else {
      for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int j = 0; j < COPY_COLS; ++j) {
          const int64_t c = c0 + j;
          if (c >= n) continue;
          if (keep(r, c)) dst[c * ldd + r] = src[c * lds + r]; else if (MODE >= 3) dst[c * ldd + r] = 0.0;
        }
      }
    }
  }
}
inline void launch_copy2d(cudaStream_t stream, double* dst, int64_t ldd, const double* src, int64_t lds, int64_t m, int64_t n, int mode) {
  const bool vec = ((((uintptr_t)dst) | ((uintptr_t)src)) & 15) == 0 && ((ldd | lds) & 1) == 0 && m >= 2;
  const int64_t work = vec ? (m + 1) / 2 : m;
  const dim3 g((unsigned)std::min<int64_t>((work + 255) / 256, 4096), (unsigned)std::min<int64_t>((n + COPY_COLS - 1) / COPY_COLS, 65535));
#define CAPK_COPY2D(M)                                                                            \
  if (vec) copy2d_kernel<M, true><<<g, 256, 0, stream>>>(dst, ldd, src, lds, m, n);               \
  else copy2d_kernel<M, false><<<g, 256, 0, stream>>>(dst, ldd, src, lds, m, n)
  switch (mode) {
    case 0: CAPK_COPY2D(0); break;
    case 1: CAPK_COPY2D(1); break;
    case 2: CAPK_COPY2D(2); break;
    case 3: CAPK_COPY2D(3); break;
    default: CAPK_COPY2D(4); break;
  }
#undef CAPK_COPY2D
}

// packed upper (reference uppertri offset X(X+1)/2+Y, src/matrix/structure.h:39) <-> full column-major
// (PACK_COLS columns per thread and pass: that many independent loads are in flight before the first store)
constexpr int PACK_COLS = 8;
__global__ void pack_upper_kernel(double* __restrict__ packed, const double* __restrict__ full, int64_t ld, int n) {
  for (int c0 = blockIdx.y * PACK_COLS; c0 < n; c0 += gridDim.y * PACK_COLS) {
    const int cmax = min(n, c0 + PACK_COLS) - 1;                     // last column of this group: rows beyond it belong to nobody
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r <= cmax; r += gridDim.x * blockDim.x) {
      double v[PACK_COLS];
#pragma unroll
      for (int j = 0; j < PACK_COLS; ++j) { const int c = c0 + j; v[j] = (c < n && r <= c) ? full[r + (int64_t)c * ld] : 0.0; }
#pragma unroll
      for (int j = 0; j < PACK_COLS; ++j) { const int c = c0 + j; if (c < n && r <= c) packed[(int64_t)c * (c + 1) / 2 + r] = v[j]; }
    }
  }
}
__global__ void unpack_upper_kernel(double* __restrict__ full, int64_t ld, const double* __restrict__ packed, int n, int zero_lower) {
  for (int c0 = blockIdx.y * PACK_COLS; c0 < n; c0 += gridDim.y * PACK_COLS) {
    const int rmax = zero_lower ? n - 1 : min(n, c0 + PACK_COLS) - 1;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r <= rmax; r += gridDim.x * blockDim.x) {
      double v[PACK_COLS];
#pragma unroll
      for (int j = 0; j < PACK_COLS; ++j) { const int c = c0 + j; v[j] = (c < n && r <= c) ? packed[(int64_t)c * (c + 1) / 2 + r] : 0.0; }
#pragma unroll
      for (int j = 0; j < PACK_COLS; ++j) { const int c = c0 + j; if (c < n && (r <= c || zero_lower)) full[r + (int64_t)c * ld] = v[j]; }
    }
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
  for (int64_t c = blockIdx.y; c < cols; c += gridDim.y)
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

// ------------------------------------------------------------------------------------------ validators
// util::residual_local (reference src/util/util.hpp:27-53): err += sum w (X - Ref)^2, ctl += sum w Ref^2 over the
// local m x n block whose element (r, c) is global (y + d r, x + d c).  Ref == nullptr: the identity (ctl counts 1 per
// element, as test/qr/validate.hpp:24-29).  upper_only: elements with global row <= column only
// (test/cholesky/validate.hpp).  w = offdiag_weight off the global diagonal (2 when only one triangle of a symmetric
// matrix was formed).  Deterministic: fixed grid, per-block partials, then one block sums them in order.
constexpr int RESID_BLOCKS = 1184, RESID_THREADS = 256;
__global__ void __launch_bounds__(RESID_THREADS) residual_partial_kernel(const double* __restrict__ X, int64_t ldx, const double* __restrict__ Ref, int64_t ldref,
                                                                          int64_t m, int64_t n, int d, int x, int y, int upper_only, double offdiag_weight,
                                                                          double* __restrict__ partial) {
  double err = 0.0, ctl = 0.0;
  const int64_t total = m * n;
  for (int64_t idx = (int64_t)blockIdx.x * RESID_THREADS + threadIdx.x; idx < total; idx += (int64_t)RESID_BLOCKS * RESID_THREADS) {
    const int64_t r = idx % m, c = idx / m, gr = y + (int64_t)d * r, gc = x + (int64_t)d * c;
    if (upper_only && gr > gc) continue;
    const double ref = Ref ? Ref[r + c * ldref] : (gr == gc ? 1.0 : 0.0);
    const double w = gr == gc ? 1.0 : offdiag_weight;
    const double diff = X[r + c * ldx] - ref;
    err += w * diff * diff;
    ctl += Ref ? w * ref * ref : w;
  }
  __shared__ double se[RESID_THREADS / 32], sc[RESID_THREADS / 32];
  for (int o = 16; o; o >>= 1) { err += __shfl_xor_sync(0xffffffffu, err, o); ctl += __shfl_xor_sync(0xffffffffu, ctl, o); }
  if ((threadIdx.x & 31) == 0) { se[threadIdx.x >> 5] = err; sc[threadIdx.x >> 5] = ctl; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double e = 0.0, cc = 0.0;
    for (int w = 0; w < RESID_THREADS / 32; ++w) { e += se[w]; cc += sc[w]; }
    partial[2 * blockIdx.x] = e; partial[2 * blockIdx.x + 1] = cc;
  }
}
__global__ void residual_final_kernel(const double* __restrict__ partial, double* __restrict__ acc) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double e = 0.0, c = 0.0;
    for (int b = 0; b < RESID_BLOCKS; ++b) { e += partial[2 * b]; c += partial[2 * b + 1]; }
    acc[0] += e; acc[1] += c;
  }
}

}  // namespace capk
