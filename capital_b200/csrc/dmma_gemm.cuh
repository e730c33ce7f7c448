// FP64 tensor-core (DMMA.8x8x4) GEMM tile engine for sm_100a.
//
// One engine serves every dense contraction on the path:
//   gemm NN/NT/TN/TT            blas::engine::_gemm   (reference src/blas/interface.hpp:43-59)
//   syrk (triangular C store)   blas::engine::_syrk   (src/blas/interface.hpp:82-97)
//   trmm (k-range skipping)     blas::engine::_trmm   (src/blas/interface.hpp:62-79)
//   split-K Gram partials       cacqr.hpp:15 (tall-skinny Q^T Q)
//
// CTA tile 128x128x16, 256 threads = 8 warps laid out 2 (M) x 4 (N); warp tile 64x32 = 8x4 DMMA
// tiles (64 accumulator doubles per thread); 4-stage cp.async (LDGSTS) pipeline, 130 KB smem,
// one CTA per SM.  DMMA issues once per 16 cycles per SM sub-partition, so issue slots are not
// the constraint; shared-memory fragment traffic is kept to one LDS.128 per two DMMA operands by
//   - K-contiguous operands: thread t of a quad takes k = 2t, 2t+1 (a permuted k order that both
//     operands share), rows XOR-swizzled so the 8 threads of an LDS.128 phase hit 8 distinct chunks;
//   - MN-contiguous operands: thread g takes rows 2g, 2g+1 (two interleaved DMMA tiles), k-rows
//     padded to 130 doubles for the same conflict-free property.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace capk {

constexpr int BM = 128, BN = 128, BK = 16;
constexpr int NTHREADS = 256;
constexpr int STAGES = 4;
constexpr int LD_MC = 130;                     // MN-contiguous tile: [k=16][130]
constexpr int TILE_DOUBLES = BK * LD_MC;       // 2080 >= 128*16 (K-contiguous tile: [mn=128][16])
constexpr int STAGE_DOUBLES = 2 * TILE_DOUBLES;
constexpr int GEMM_SMEM_BYTES = STAGES * STAGE_DOUBLES * 8;  // 133120

enum : int { KR_NONE = 0, KR_TILE_M = 1, KR_TILE_N = 2 };   // k-range bound follows the m- or n-tile
enum : int { CTRI_FULL = 0, CTRI_UPPER = 1, CTRI_LOWER = 2 };

struct GemmParams {
  const double* A; const double* B; double* C; const double* Cin;
  int64_t lda, ldb, ldc, ldcin;
  int M, N, K;
  double alpha, beta;
  int klo_mode, khi_mode;   // k_begin = tile*128 / k_end = (tile+1)*128 for triangular operands
  int c_tri;                // store only one triangle of C (and skip tiles outside it)
  int ksplit;               // >1: blockIdx.y selects a k-slab, raw partial written at C + y*c_split_stride
  int k_per_split;
  int64_t c_split_stride;
  int tiles_m, tiles_n;
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// Copy one 2-double chunk global->shared with bounds: `contig_left` = elements still valid along the
// contiguous direction (<=0 none), `other_ok` = the strided index is in range.  Out-of-range bytes are
// zero-filled by cp.async (src-size < cp-size), so edge tiles need no separate path.
template <int VEC>
__device__ __forceinline__ void load_chunk(double* dst, const double* base, const double* src, int contig_left, bool other_ok) {
  int valid = other_ok ? (contig_left < 0 ? 0 : (contig_left > 2 ? 2 : contig_left)) : 0;
  if (VEC == 2) {
    cp_async16(dst, valid ? src : base, valid * 8);
  } else {
    cp_async8(dst, valid > 0 ? src : base, valid > 0 ? 8 : 0);
    cp_async8(dst + 1, valid > 1 ? src + 1 : base, valid > 1 ? 8 : 0);
  }
}

// K-contiguous operand: element (mn, k) at ptr[mn*ld + k].  Shared tile [128][16], 16-byte chunk c of
// row r stored at chunk c ^ ((r&1)<<2).
template <int VEC>
__device__ __forceinline__ void load_tile_kc(double* s, const double* ptr, int64_t ld, int mn0, int mn_lim, int k0, int k_lim, int tid) {
  const int c = tid & 7;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = (tid >> 3) + 32 * i;
    const int mn = mn0 + r, k = k0 + 2 * c;
    double* dst = s + r * 16 + ((c ^ ((r & 1) << 2)) << 1);
    load_chunk<VEC>(dst, ptr, ptr + (int64_t)mn * ld + k, k_lim - k, mn < mn_lim);
  }
}

// MN-contiguous operand: element (mn, k) at ptr[k*ld + mn].  Shared tile [16][130].
template <int VEC>
__device__ __forceinline__ void load_tile_mc(double* s, const double* ptr, int64_t ld, int mn0, int mn_lim, int k0, int k_lim, int tid) {
  const int c = tid & 63;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int kr = (tid >> 6) + 4 * i;
    const int mn = mn0 + 2 * c, k = k0 + kr;
    double* dst = s + kr * LD_MC + 2 * c;
    load_chunk<VEC>(dst, ptr, ptr + (int64_t)k * ld + mn, mn_lim - mn, k < k_lim);
  }
}

// 16 k-columns of the 128x128 tile: 2 halves x (fragment loads, then 2 x 32 DMMA).
template <bool TA, bool TB>
__device__ __forceinline__ void compute_stage(const double* __restrict__ sA, const double* __restrict__ sB,
                                              double (&acc)[8][4][2], int wm, int wn, int g, int t) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    double a[2][8], b[2][4];
    if (TA) {  // op(A) K-contiguous
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = wm * 64 + i * 8 + g;
        const double2 v = *reinterpret_cast<const double2*>(sA + r * 16 + (((h * 4 + t) ^ ((g & 1) << 2)) << 1));
        a[0][i] = v.x; a[1][i] = v.y;
      }
    } else {   // op(A) M-contiguous: tile 2j+e holds rows 16j + 2g + e
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const double2 v = *reinterpret_cast<const double2*>(sA + (8 * h + 2 * t + e) * LD_MC + wm * 64 + 16 * j + 2 * g);
          a[e][2 * j] = v.x; a[e][2 * j + 1] = v.y;
        }
    }
    if (!TB) {  // op(B) K-contiguous
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = wn * 32 + j * 8 + g;
        const double2 v = *reinterpret_cast<const double2*>(sB + r * 16 + (((h * 4 + t) ^ ((g & 1) << 2)) << 1));
        b[0][j] = v.x; b[1][j] = v.y;
      }
    } else {    // op(B) N-contiguous
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const double2 v = *reinterpret_cast<const double2*>(sB + (8 * h + 2 * t + e) * LD_MC + wn * 32 + 16 * j + 2 * g);
          b[e][2 * j] = v.x; b[e][2 * j + 1] = v.y;
        }
    }
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(acc[i][j], a[e][i], b[e][j]);
  }
}

// Row / column of accumulator (tile index i or j, position p inside the 8-wide DMMA tile).
template <bool MC>
__device__ __forceinline__ int frag_pos(int tile, int p) { return MC ? 16 * (tile >> 1) + 2 * p + (tile & 1) : tile * 8 + p; }

// Compute and store one 128x128 tile of C.  All 256 threads participate.
template <bool TA, bool TB, int VEC>
__device__ __forceinline__ void gemm_tile(const GemmParams& p, double* smem, int tm, int tn, int kz) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wm = warp >> 2, wn = warp & 3, g = lane >> 2, t = lane & 3;
  const int m0 = tm * BM, n0 = tn * BN;

  int k_begin = 0, k_end = p.K;
  if (p.klo_mode == KR_TILE_M) k_begin = m0; else if (p.klo_mode == KR_TILE_N) k_begin = n0;
  if (p.khi_mode == KR_TILE_M) k_end = min(p.K, m0 + BM); else if (p.khi_mode == KR_TILE_N) k_end = min(p.K, n0 + BN);
  if (p.ksplit > 1) { k_begin = max(k_begin, kz * p.k_per_split); k_end = min(k_end, (kz + 1) * p.k_per_split); }
  const int nk = k_end > k_begin ? (k_end - k_begin + BK - 1) / BK : 0;

  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  auto load_stage = [&](int stage, int kt) {
    double* sA = smem + stage * STAGE_DOUBLES;
    double* sB = sA + TILE_DOUBLES;
    const int k0 = k_begin + kt * BK;
    if (TA) load_tile_kc<VEC>(sA, p.A, p.lda, m0, p.M, k0, k_end, tid); else load_tile_mc<VEC>(sA, p.A, p.lda, m0, p.M, k0, k_end, tid);
    if (!TB) load_tile_kc<VEC>(sB, p.B, p.ldb, n0, p.N, k0, k_end, tid); else load_tile_mc<VEC>(sB, p.B, p.ldb, n0, p.N, k0, k_end, tid);
  };

  __syncthreads();  // the previous tile's consumers are done with shared memory
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < nk) load_stage(s, s);
    cp_async_commit();
  }
  for (int kt = 0; kt < nk; ++kt) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    const int nxt = kt + STAGES - 1;
    if (nxt < nk) load_stage(nxt % STAGES, nxt);
    cp_async_commit();
    const double* sA = smem + (kt % STAGES) * STAGE_DOUBLES;
    compute_stage<TA, TB>(sA, sA + TILE_DOUBLES, acc, wm, wn, g, t);
  }
  cp_async_wait<0>();

  // epilogue: C = alpha*acc + beta*Cin (raw partial when split-K)
  double* C = p.C;
  double alpha = p.alpha, beta = p.beta;
  const double* Cin = p.Cin;
  if (p.ksplit > 1) { C += (int64_t)kz * p.c_split_stride; alpha = 1.0; beta = 0.0; }
  // batches of two row tiles: every read of Cin is in flight before the first write (C may alias Cin, so element-by-element
  // code makes each load wait for the store before it -- 64 dependent L2 round trips per thread)
#pragma unroll
  for (int i0 = 0; i0 < 8; i0 += 2) {
    double o[2][4][2];
    uint32_t live = 0;
#pragma unroll
    for (int ii = 0; ii < 2; ++ii) {
      const int r = m0 + wm * 64 + frag_pos<!TA>(i0 + ii, g);
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int c = n0 + wn * 32 + frag_pos<TB>(j, 2 * t + e);
          bool ok = (r < p.M) && (c < p.N);
          if (p.c_tri == CTRI_UPPER && r > c) ok = false;
          if (p.c_tri == CTRI_LOWER && r < c) ok = false;
          live |= (ok ? 1u : 0u) << (ii * 8 + j * 2 + e);
          o[ii][j][e] = (ok && beta != 0.0) ? Cin[r + (int64_t)c * p.ldcin] : 0.0;
        }
    }
#pragma unroll
    for (int ii = 0; ii < 2; ++ii) {
      const int r = m0 + wm * 64 + frag_pos<!TA>(i0 + ii, g);
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int c = n0 + wn * 32 + frag_pos<TB>(j, 2 * t + e);
          if ((live >> (ii * 8 + j * 2 + e)) & 1u) C[r + (int64_t)c * p.ldc] = fma(beta, o[ii][j][e], alpha * acc[i0 + ii][j][e]);
        }
    }
  }
}

// Grid: x = tile index (grouped raster for L2 reuse; triangular enumeration when c_tri on a square C),
//       y = k-slab (split-K).
template <bool TA, bool TB, int VEC>
__global__ void __launch_bounds__(NTHREADS, 1) dmma_gemm_kernel(const GemmParams p) {
  extern __shared__ __align__(16) double smem[];
  int tm, tn;
  const int bid = blockIdx.x;
  if (p.c_tri != CTRI_FULL && p.tiles_m == p.tiles_n) {
    // bid enumerates the T(T+1)/2 tiles of one triangle, column by column
    int col = 0, rem = bid;
    while (rem > col) { rem -= col + 1; ++col; }
    tm = (p.c_tri == CTRI_UPPER) ? rem : col;
    tn = (p.c_tri == CTRI_UPPER) ? col : rem;
  } else {
    constexpr int GROUP = 8;
    const int per_group = GROUP * p.tiles_n;
    const int first_m = (bid / per_group) * GROUP;
    const int gsize = min(p.tiles_m - first_m, GROUP);
    tm = first_m + (bid % per_group) % gsize;
    tn = (bid % per_group) / gsize;
    if (p.c_tri == CTRI_UPPER && tn * BN + BN - 1 < tm * BM) return;
    if (p.c_tri == CTRI_LOWER && tm * BM + BM - 1 < tn * BN) return;
  }
  gemm_tile<TA, TB, VEC>(p, smem, tm, tn, blockIdx.y);
}

// In-place panel variant for B <- B*op(T) (trmm Right): one CTA owns a 128-row block of B (= both the
// A operand and C) and walks the column tiles in the order that never overwrites a still-needed input:
// right-to-left when only k <= tile_n contributes (Upper/NoTrans, Lower/Trans), left-to-right otherwise.
template <bool TB, int VEC>
__global__ void __launch_bounds__(NTHREADS, 1) dmma_trmm_right_inplace_kernel(const GemmParams p) {
  extern __shared__ __align__(16) double smem[];
  const bool descending = (p.khi_mode == KR_TILE_N);
  for (int tm = blockIdx.x; tm < p.tiles_m; tm += gridDim.x)
    for (int s = 0; s < p.tiles_n; ++s) {
      const int tn = descending ? p.tiles_n - 1 - s : s;
      gemm_tile<false, TB, VEC>(p, smem, tm, tn, 0);
    }
}

// ------------------------------------------------------------------------------------------ small products
// Products whose C has at most a few 128x128 tiles (the inner levels of cholinv / trtri) would sit on one or two
// SMs with the big-tile engines: one 128^3 tile is 17 us at a single SM's DMMA rate, and these products are what
// the recursion spends its time in (55 % of cholinv(2048), profiles/r01_launches_cholinv2048.csv).
// Here one CTA owns a 32x32 block of C and its 8 warps split the k-range: every warp issues ALL its DMMA
// fragment loads up front straight from global memory (operands are L2-resident; a fragment load uses whole
// 32-byte sectors in either storage order), runs its DMMAs, and the partial blocks are summed through shared
// memory in a fixed order (deterministic) and stored with coalesced writes.  A 128^3 product = 16 CTAs x 8 warps
// with 4 k-steps each.  k-ranges of triangular operands are applied per 32-block.  Same GemmParams, no split-K.
constexpr int SMALL_MAX_MN = 512;                              // largest M, N routed here (K <= 512)
constexpr int SMALL_THREADS = 256;
constexpr int SMALL_BATCH = 8;                                 // k-steps (of 4) in flight per warp
constexpr int SMALL_SMEM_BYTES = 8 * 32 * 32 * 8;              // 8 warps x 32 accumulators x 32 lanes

template <bool TA, bool TB>
__global__ void __launch_bounds__(SMALL_THREADS) dmma_small_gemm_kernel(const GemmParams p) {
  extern __shared__ __align__(16) double red[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int tiles_m = (p.M + 31) / 32;
  const int r0 = (blockIdx.x % tiles_m) * 32, c0 = (blockIdx.x / tiles_m) * 32;
  if (p.c_tri == CTRI_UPPER && c0 + 31 < r0) return;
  if (p.c_tri == CTRI_LOWER && r0 + 31 < c0) return;
  int k_begin = 0, k_end = p.K;
  if (p.klo_mode == KR_TILE_M) k_begin = r0; else if (p.klo_mode == KR_TILE_N) k_begin = c0;
  if (p.khi_mode == KR_TILE_M) k_end = min(p.K, r0 + 32); else if (p.khi_mode == KR_TILE_N) k_end = min(p.K, c0 + 32);
  const int nsteps = max(0, (k_end - k_begin + 3) / 4), per = (nsteps + 7) / 8;
  const int s_begin = warp * per, s_end = min(nsteps, s_begin + per);
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  const double* pa[4]; const double* pb[4]; bool oka[4], okb[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + 8 * i + g, c = c0 + 8 * i + g;
    oka[i] = r < p.M; okb[i] = c < p.N;
    pa[i] = TA ? p.A + (int64_t)r * p.lda : p.A + r;           // + k (TA) or + k * lda
    pb[i] = TB ? p.B + c : p.B + (int64_t)c * p.ldb;           // + k * ldb (TB) or + k
  }
  for (int sb = s_begin; sb < s_end; sb += SMALL_BATCH) {
    double af[SMALL_BATCH][4], bf[SMALL_BATCH][4];
#pragma unroll
    for (int u = 0; u < SMALL_BATCH; ++u) {
      const int k = k_begin + 4 * (sb + u) + t;
      const bool kok = (sb + u < s_end) && k < k_end;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        af[u][i] = (kok && oka[i]) ? __ldg(TA ? pa[i] + k : pa[i] + (int64_t)k * p.lda) : 0.0;
        bf[u][i] = (kok && okb[i]) ? __ldg(TB ? pb[i] + (int64_t)k * p.ldb : pb[i] + k) : 0.0;
      }
    }
#pragma unroll
    for (int u = 0; u < SMALL_BATCH; ++u) {
      if (sb + u < s_end) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma884(acc[i][j], af[u][i], bf[u][j]);
      }
    }
  }
  // partial blocks -> shared memory [warp][accumulator][lane]
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) red[(warp * 32 + (i * 4 + j) * 2 + e) * 32 + lane] = acc[i][j][e];
  __syncthreads();
  // thread -> row (tid & 31) and four columns: fixed-order sum over the warps, coalesced store
  const int rr = threadIdx.x & 31, r = r0 + rr;
  if (r >= p.M) return;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int cc = (threadIdx.x >> 5) + 8 * q, c = c0 + cc;
    if (c >= p.N) continue;
    if (p.c_tri == CTRI_UPPER && r > c) continue;
    if (p.c_tri == CTRI_LOWER && r < c) continue;
    const int slot = (((rr >> 3) * 4 + (cc >> 3)) * 2 + (cc & 1)) * 32 + (rr & 7) * 4 + ((cc & 7) >> 1);
    double sum = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) sum += red[w * 1024 + slot];
    double v = p.alpha * sum;
    if (p.beta != 0.0) v += p.beta * p.Cin[r + (int64_t)c * p.ldcin];
    p.C[r + (int64_t)c * p.ldc] = v;
  }
}

// Deterministic split-K reduction: C = alpha * sum_z part[z] + beta*C over one triangle or the full matrix.
__global__ void splitk_reduce_kernel(const double* __restrict__ part, int64_t part_stride, int ld_part, int ksplit,
                                     double* C, int64_t ldc, int M, int N, double alpha, double beta, int c_tri) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= M) return;
  for (int c = blockIdx.y; c < N; c += gridDim.y) {      // (columns strided: any N fits the 65535 limit of gridDim.y)
    if (c_tri == CTRI_UPPER && r > c) continue;
    if (c_tri == CTRI_LOWER && r < c) continue;
    double s = 0.0;
    for (int z = 0; z < ksplit; ++z) s += part[(int64_t)z * part_stride + r + (int64_t)c * ld_part];
    double v = alpha * s;
    if (beta != 0.0) v += beta * C[r + (int64_t)c * ldc];
    C[r + (int64_t)c * ldc] = v;
  }
}

}  // namespace capk
