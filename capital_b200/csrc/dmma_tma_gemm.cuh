// Warp-specialised, persistent FP64 DMMA tile engine for sm_100a: TMA (cp.async.bulk.tensor) producer
// warp + mbarrier ring + 8 consumer warps that do nothing but LDS.128 and DMMA.8x8x4.
//
// Same contractions as dmma_gemm.cuh (gemm / syrk / trmm / split-K Gram; reference
// src/blas/interface.hpp:43-97), used whenever the operands satisfy TMA's alignment rules (16-byte
// base, even leading dimension); the cp.async engine stays as the path for everything else.
//
// Why a second engine (profiles/r01_ncu_full_cpasync_engine.txt): in the cp.async kernel every warp
// also computes bounds/addresses and meets the others at a CTA barrier once per 16 k-columns; the DMMA
// pipe was active 80-84 % of the time.  Here
//   * one elected producer lane issues the TMA boxes of a stage (out-of-range parts are zero-filled by
//     the hardware, so there is no bounds arithmetic at all) and runs ahead across tiles;
//   * consumers wait on a "full" mbarrier, read fragments, issue DMMAs, release the stage through an
//     "empty" mbarrier -- no CTA-wide barrier anywhere in the main loop, the two warps that share an
//     SM sub-partition drift out of phase and cover each other's fragment-load latency;
//   * CTAs are persistent (grid = #SMs), so the epilogue of one tile overlaps the loads of the next;
//   * tiles that touch the diagonal of a triangular operand or of a triangular C skip, at 16x16-block
//     granularity, the DMMAs that would only multiply structural zeros / compute unwanted entries.
//     Column pairs are dealt to warps as {wn, 7-wn} so that the skipped work is balanced over the four
//     SM sub-partitions (each keeps 9/16 of a diagonal block's work).
//
// Shared-memory tiles use the TMA 128-byte swizzle (16-byte chunk index XOR (row & 7)):
//   K-contiguous operand : one box  [128 mn][16 k]       row = 128 B
//   MN-contiguous operand: 8 boxes  [16 k][16 mn] (2 KB) row = 128 B
// and the fragment <-> row assignment below makes every LDS.128 phase hit 8 distinct chunks.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "dmma_gemm.cuh"

namespace capk {

constexpr int T_STAGES = 6;
constexpr int T_TILE_BYTES = 128 * 16 * 8;            // 16 KB per operand per stage
constexpr int T_STAGE_BYTES = 2 * T_TILE_BYTES;       // 32 KB
constexpr int T_CONSUMER_WARPS = 8;
constexpr int T_THREADS = (T_CONSUMER_WARPS + 4) * 32;  // + one producer warpgroup (register budgets move in units of 4 warps)
constexpr int T_REGS_CONSUMER = 240, T_REGS_PRODUCER = 24;
constexpr int T_SMEM_BYTES = T_STAGES * T_STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*mbarriers*/;

enum : int { SCHED_RASTER = 0, SCHED_TRI = 1, SCHED_ROWPANEL_DESC = 2, SCHED_ROWPANEL_ASC = 3, SCHED_TM_DESC = 4, SCHED_STAIR = 5 };
// which operand is triangular on the diagonal k-block (the relation that holds for NON-zero entries)
enum : int { DIAG_NONE = 0, DIAG_A_LE = 1 /* r <= k */, DIAG_A_GE = 2 /* r >= k */, DIAG_B_LE = 3 /* k <= c */, DIAG_B_GE = 4 /* k >= c */ };

struct TmaGemmParams {
  double* C; const double* Cin;
  int64_t ldc, ldcin;
  int M, N, K;
  double alpha, beta;
  int klo_mode, khi_mode;
  int c_tri;
  int diag;                 // DIAG_*: skip structural zeros of the triangular operand
  int ksplit, k_per_split;
  int64_t c_split_stride;
  int tiles_m, tiles_n;
  int sched;
  int units, subs;          // persistent loop: unit = blockIdx.x + i*gridDim.x, `subs` tiles per unit
  int c_vec2;               // C/Cin 16-byte aligned with even ld: paired-row stores allowed
  // SCHED_STAIR: C is a run of block columns of a 1-D block-cyclic (boustrophedon) matrix, rows start at global block
  // stair_row0; only tiles on or above the GLOBAL diagonal exist (one launch = the whole trailing update of a rank)
  int stair_T, stair_P, stair_me, stair_r_first, stair_row0;   // tiles per block, ranks, this rank, first local block, first row block
};

// ------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------------------------------ work decode
// stage kt of a tile covers k0 = k_first + kt*k_step .. k0+15 (k_step = +-16: row-panel schedules walk k in
// alternating directions so that a panel block is re-read while it is still in L2); the diagonal 128-block
// of a triangular operand starts at k = diag_origin.
struct TileInfo { int tm, tn, kz, k_first, k_step, nk, diag_origin; bool tri_diag; };
struct TrueTag { static constexpr bool value = true; };
struct FalseTag { static constexpr bool value = false; };

// SCHED_STAIR: column-major walk over the tiles of every block column that lie on or above the global diagonal.  The column
// tile `ct` of local block rb (global block g under the boustrophedon deal) has (g - row0) * T + ct + 1 row tiles, the last of
// which is triangular.  Host-callable: tests/test_potrf1d_host.py enumerates it on the CPU through cap_debug_stair_decode.
__host__ __device__ inline void stair_decode(int T, int P, int me, int r_first, int row0, int tiles_n, int unit, int& tm, int& tn, bool& tri) {
  int rem = unit;
  for (int bi = 0;; ++bi) {
    const int rb = r_first + bi;
    const int g = rb * P + ((rb & 1) ? P - 1 - me : me);
    const int Tb = (T < tiles_n - bi * T) ? T : tiles_n - bi * T;
    const int base = (g - row0) * T + 1;
    const int cnt = Tb * base + Tb * (Tb - 1) / 2;
    if (rem < cnt || (bi + 1) * T >= tiles_n) {
      int ct = 0;
      while (ct + 1 < Tb && rem >= base + ct) { rem -= base + ct; ++ct; }
      tn = bi * T + ct; tm = rem; tri = (rem == base + ct - 1);
      return;
    }
    rem -= cnt;
  }
}

__device__ __forceinline__ TileInfo decode_tile(const TmaGemmParams& p, int unit, int sub) {
  TileInfo t;
  int tile = unit;
  t.kz = 0;
  t.tri_diag = false;
  if (p.sched == SCHED_STAIR) {
    stair_decode(p.stair_T, p.stair_P, p.stair_me, p.stair_r_first, p.stair_row0, p.tiles_n, unit, t.tm, t.tn, t.tri_diag);
  } else
  if (p.sched == SCHED_ROWPANEL_DESC) { t.tm = unit; t.tn = p.tiles_n - 1 - sub; }
  else if (p.sched == SCHED_ROWPANEL_ASC) { t.tm = unit; t.tn = sub; }
  else {
    const int per_k = (p.sched == SCHED_TRI) ? p.tiles_m * (p.tiles_m + 1) / 2 : p.tiles_m * p.tiles_n;
    if (p.ksplit > 1) { t.kz = unit / per_k; tile = unit - t.kz * per_k; }
    if (p.sched == SCHED_TRI) {
      int col = 0, rem = tile;
      while (rem > col) { rem -= col + 1; ++col; }
      t.tm = (p.c_tri == CTRI_UPPER) ? rem : col;
      t.tn = (p.c_tri == CTRI_UPPER) ? col : rem;
    } else if (p.sched == SCHED_TM_DESC) {   // longest-k tiles first when k grows with tm
      t.tm = p.tiles_m - 1 - tile / p.tiles_n; t.tn = tile % p.tiles_n;
    } else if (p.klo_mode == KR_TILE_M) {    // k shrinks with tm: plain tm-major order is already longest-first
      t.tm = tile / p.tiles_n; t.tn = tile % p.tiles_n;
    } else {
      constexpr int GROUP = 8;
      const int per_group = GROUP * p.tiles_n;
      const int first_m = (tile / per_group) * GROUP;
      const int gsize = min(p.tiles_m - first_m, GROUP);
      t.tm = first_m + (tile % per_group) % gsize;
      t.tn = (tile % per_group) / gsize;
    }
  }
  const int m0 = t.tm * BM, n0 = t.tn * BN;
  int k_begin = 0, k_end = p.K;
  if (p.klo_mode == KR_TILE_M) k_begin = m0; else if (p.klo_mode == KR_TILE_N) k_begin = n0;
  if (p.khi_mode == KR_TILE_M) k_end = min(p.K, m0 + BM); else if (p.khi_mode == KR_TILE_N) k_end = min(p.K, n0 + BN);
  if (p.ksplit > 1) { k_begin = max(k_begin, t.kz * p.k_per_split); k_end = min(k_end, (t.kz + 1) * p.k_per_split); }
  t.nk = k_end > k_begin ? (k_end - k_begin + BK - 1) / BK : 0;
  const bool descending = (p.sched == SCHED_ROWPANEL_DESC || p.sched == SCHED_ROWPANEL_ASC) && (sub & 1);
  t.k_step = descending ? -BK : BK;
  t.k_first = descending ? k_begin + (t.nk - 1) * BK : k_begin;
  t.diag_origin = (p.diag == DIAG_A_LE || p.diag == DIAG_A_GE) ? m0 : n0;
  if (p.sched != SCHED_STAIR) t.tri_diag = (p.c_tri != CTRI_FULL && t.tm == t.tn);
  return t;
}

// skip tiles entirely outside the stored triangle of C (only reachable with the rectangular raster)
__device__ __forceinline__ bool tile_outside_tri(const TmaGemmParams& p, const TileInfo& t) {
  if (p.sched == SCHED_STAIR) return false;
  if (p.c_tri == CTRI_UPPER && t.tn * BN + BN - 1 < t.tm * BM) return true;
  if (p.c_tri == CTRI_LOWER && t.tm * BM + BM - 1 < t.tn * BN) return true;
  return false;
}

// ------------------------------------------------------------------------------------------ consumer side
// Lane (g = lane>>2, t = lane&3).  pi(g) spreads the 8 DMMA rows of a K-contiguous tile over smem rows
// whose (row & 7) differ in bit 2 between neighbouring g, which is what makes the swizzled LDS.128
// conflict-free (lanes 0..7 of a phase are g in {2q, 2q+1} x t in 0..3).
__device__ __forceinline__ int pi8(int g) { return ((g & 1) << 2) | (g >> 1); }

// local row of accumulator row-tile i (0..7) for this lane's g
template <bool A_KC>
__device__ __forceinline__ int acc_row(int wm, int i, int g) {
  return A_KC ? wm * 64 + 8 * i + pi8(g) : wm * 64 + 16 * (i >> 1) + 2 * g + (i & 1);
}
// local column of accumulator col-tile jj (0..3), C-fragment column index cidx (0..7)
template <bool B_KC>
__device__ __forceinline__ int acc_col(int wn, int jj, int cidx) {
  const int pair = (jj < 2) ? wn : 7 - wn;
  return B_KC ? 16 * pair + 8 * (jj & 1) + pi8(cidx) : 16 * pair + 2 * cidx + (jj & 1);
}

// 8-bit block mask, bit (j*2+q): row group j (16 rows: wm*64+16j..) x column pair q (16 cols: pair wn or 7-wn)
__device__ __forceinline__ uint32_t diag_mask(int kind, int s, int wm, int wn) {
  uint32_t m = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int rg = 4 * wm + j, pr = q ? 7 - wn : wn;
      bool need = true;
      if (kind == DIAG_A_LE) need = rg <= s;
      else if (kind == DIAG_A_GE) need = rg >= s;
      else if (kind == DIAG_B_LE) need = s <= pr;
      else if (kind == DIAG_B_GE) need = s >= pr;
      m |= (need ? 1u : 0u) << (j * 2 + q);
    }
  return m;
}
__device__ __forceinline__ uint32_t ctri_mask(int c_tri, int wm, int wn) {
  uint32_t m = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int rg = 4 * wm + j, pr = q ? 7 - wn : wn;
      const bool need = (c_tri == CTRI_UPPER) ? rg <= pr : rg >= pr;
      m |= (need ? 1u : 0u) << (j * 2 + q);
    }
  return m;
}

// Per-thread fragment addressing, computed once per kernel: every LDS.128 of a stage is
// [stage base + one of these + compile-time immediate], so the main loop carries no address arithmetic.
//   K-contiguous A : row r = wm*64 + 8i + pi(g), chunk (4h+t)^pi(g)         -> a0 ^ (h*64) + i*1024
//   MN-contiguous A: box 4wm+j, k-row 8h+2t+e, chunk g^(2t+e)                -> a_e + h*1024 + j*2048
//   K-contiguous B : row 16*pair_q + 8y + pi(g), chunk (4h+t)^pi(g)          -> b_q ^ (h*64) + y*1024
//   MN-contiguous B: box pair_q, k-row 8h+2t+e, chunk g^(2t+e)               -> b_{q,e} + h*1024
struct FragAddr { uint32_t a[2]; uint32_t b[4]; };

template <bool A_KC, bool B_KC>
__device__ __forceinline__ FragAddr frag_addr(int wm, int wn, int g, int t) {
  FragAddr f;
  if (A_KC) { f.a[0] = (uint32_t)((wm * 64 + pi8(g)) * 128 + ((t ^ pi8(g)) << 4)); f.a[1] = f.a[0] ^ 64u; }
  else {
#pragma unroll
    for (int e = 0; e < 2; ++e) f.a[e] = (uint32_t)(wm * 8192 + (2 * t + e) * 128 + ((g ^ (2 * t + e)) << 4));
  }
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int pair = q ? 7 - wn : wn;
    if (B_KC) { f.b[2 * q] = (uint32_t)(T_TILE_BYTES + (16 * pair + pi8(g)) * 128 + ((t ^ pi8(g)) << 4)); f.b[2 * q + 1] = f.b[2 * q] ^ 64u; }
    else {
#pragma unroll
      for (int e = 0; e < 2; ++e) f.b[2 * q + e] = (uint32_t)(T_TILE_BYTES + pair * 2048 + (2 * t + e) * 128 + ((g ^ (2 * t + e)) << 4));
    }
  }
  return f;
}

template <int IMM>
__device__ __forceinline__ void lds128(double& x, double& y, uint32_t addr) {
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(x), "=d"(y) : "r"(addr), "n"(IMM));
}

// Fragments of one half stage (8 k-columns = 2 DMMA k-steps e = 0, 1), split so that the software pipeline holds
// at most 32 doubles of fragments (B of the half stage + A of one 32-row half, each double-buffered): with whole
// half stages double-buffered (48 doubles) the consumer loop spilled ~18 registers to local memory and paid for
// them in long-scoreboard stalls at every stage boundary (profiles/r01_ncu_full_cqr2_kernels_4194304x1024.txt).
struct FragA { double v[2][4]; };   // [e][row tile 4R + i]
struct FragB { double v[2][4]; };   // [e][column tile]

// sbase: shared-space address of the stage (A tile, then B tile at +16 KB); H: which half of its 16 k-columns;
// R: which 32-row half of the warp's 64 rows.
template <bool A_KC, int H, int R>
__device__ __forceinline__ void load_a(uint32_t sbase, const FragAddr& f, FragA& fr) {
  if (A_KC) {   // v[0][i] = k 8h+2t, v[1][i] = k 8h+2t+1 of row tile 4R+i
    const uint32_t aa = sbase + f.a[H];
    lds128<(4 * R + 0) * 1024>(fr.v[0][0], fr.v[1][0], aa); lds128<(4 * R + 1) * 1024>(fr.v[0][1], fr.v[1][1], aa);
    lds128<(4 * R + 2) * 1024>(fr.v[0][2], fr.v[1][2], aa); lds128<(4 * R + 3) * 1024>(fr.v[0][3], fr.v[1][3], aa);
  } else {      // v[e][2jl], v[e][2jl+1] = rows 16j+2g, 16j+2g+1 (j = 2R + jl) at k 8h+2t+e
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const uint32_t aa = sbase + f.a[e];
      lds128<H * 1024 + (2 * R + 0) * 2048>(fr.v[e][0], fr.v[e][1], aa);
      lds128<H * 1024 + (2 * R + 1) * 2048>(fr.v[e][2], fr.v[e][3], aa);
    }
  }
}
template <bool B_KC, int H>
__device__ __forceinline__ void load_b(uint32_t sbase, const FragAddr& f, FragB& fr) {
  if (B_KC) {   // v[0][2q+y], v[1][2q+y]: column tile y of pair q
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const uint32_t bb = sbase + f.b[2 * q + H];
      lds128<0>(fr.v[0][2 * q], fr.v[1][2 * q], bb);
      lds128<1024>(fr.v[0][2 * q + 1], fr.v[1][2 * q + 1], bb);
    }
  } else {      // v[e][2q], v[e][2q+1] = columns 16*pair+2g, +1 at k 8h+2t+e
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (int e = 0; e < 2; ++e) lds128<H * 1024>(fr.v[e][2 * q], fr.v[e][2 * q + 1], sbase + f.b[2 * q + e]);
  }
}

// One k-step of a 16x16 block (2x2 DMMA tiles) behind a real, warp-uniform branch.  Written as a single
// asm statement because the compiler otherwise predicates the four DMMAs, and a predicated-off DMMA
// still costs its issue slot and the fixed spacing to the next one (profiles/r01_ncu_tma_engine.txt).
__device__ __forceinline__ void dmma_block4(double (&c00)[2], double (&c01)[2], double (&c10)[2], double (&c11)[2],
                                            double a0, double a1, double b0, double b1, uint32_t on) {
  // (a loop that runs `on` = 0 or 1 times: ptxas if-converts a plain forward branch over four DMMAs back
  //  into predication, but it cannot predicate a back edge)
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .u32 cnt;\n"
      "mov.u32 cnt, %12;\n"
      "LOOP_%=:\n"
      "setp.eq.u32 p, cnt, 0;\n"
      "@p bra.uni SKIP_%=;\n"
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%8}, {%10}, {%0,%1};\n"
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%2,%3}, {%8}, {%11}, {%2,%3};\n"
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%4,%5}, {%9}, {%10}, {%4,%5};\n"
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%6,%7}, {%9}, {%11}, {%6,%7};\n"
      "sub.u32 cnt, cnt, 1;\n"
      "bra.uni LOOP_%=;\n"
      "SKIP_%=:\n"
      "}\n"
      : "+d"(c00[0]), "+d"(c00[1]), "+d"(c01[0]), "+d"(c01[1]), "+d"(c10[0]), "+d"(c10[1]), "+d"(c11[0]), "+d"(c11[1])
      : "d"(a0), "d"(a1), "d"(b0), "d"(b1), "r"(on));
}

// The DMMAs of one half stage on the 32-row half R of the warp's 64 x 32 accumulator block.
template <int R>
__device__ __forceinline__ void mma_full(double (&acc)[8][4][2], const FragA& fa, const FragB& fb) {
#pragma unroll
  for (int e = 0; e < 2; ++e)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) dmma884(acc[4 * R + i][jj], fa.v[e][i], fb.v[e][jj]);
}
// Masked variant, only used by the separate loop over the k-stages that touch a diagonal block (and by triangular
// C tiles).  mask bit (j*2+q): row group j (16 rows) x column pair q.  Triangular operands produce one of a few
// shapes per half -- everything, one row group, one column pair, nothing -- each with straight-line code behind a
// warp-uniform branch; anything else goes block by block.
template <int R>
__device__ __forceinline__ void mma_masked(double (&acc)[8][4][2], const FragA& fa, const FragB& fb, uint32_t mask) {
  const uint32_t m4 = (mask >> (4 * R)) & 0xFu;
  if (m4 == 0xFu) {
    mma_full<R>(acc, fa, fb);
  } else if (m4 == 0x3u) {                        // first row group, both column pairs
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) dmma884(acc[4 * R + i][jj], fa.v[e][i], fb.v[e][jj]);
  } else if (m4 == 0xCu) {                        // second row group
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int i = 2; i < 4; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) dmma884(acc[4 * R + i][jj], fa.v[e][i], fb.v[e][jj]);
  } else if (m4 == 0x5u) {                        // column pair 0, both row groups
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) dmma884(acc[4 * R + i][jj], fa.v[e][i], fb.v[e][jj]);
  } else if (m4 == 0xAu) {                        // column pair 1
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jj = 2; jj < 4; ++jj) dmma884(acc[4 * R + i][jj], fa.v[e][i], fb.v[e][jj]);
  } else if (m4 != 0u) {
    // e outermost: the two updates of an accumulator are as far apart as the active blocks allow
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int jl = 0; jl < 2; ++jl)
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int j = 2 * R + jl;
          dmma_block4(acc[2 * j][2 * q], acc[2 * j][2 * q + 1], acc[2 * j + 1][2 * q], acc[2 * j + 1][2 * q + 1],
                      fa.v[e][2 * jl], fa.v[e][2 * jl + 1], fb.v[e][2 * q], fb.v[e][2 * q + 1], (mask >> (j * 2 + q)) & 1u);
        }
  }
}
template <bool MASKED, int R>
__device__ __forceinline__ void mma_part(double (&acc)[8][4][2], const FragA& fa, const FragB& fb, uint32_t mask) {
  if (MASKED) mma_masked<R>(acc, fa, fb, mask); else mma_full<R>(acc, fa, fb);
}

template <bool A_KC, bool B_KC>
__device__ __forceinline__ void store_tile(const TmaGemmParams& p, const TileInfo& ti, double (&acc)[8][4][2], int wm, int wn, int g, int t) {
  double* C = p.C;
  double alpha = p.alpha, beta = p.beta;
  if (p.ksplit > 1) { C += (int64_t)ti.kz * p.c_split_stride; alpha = 1.0; beta = 0.0; }
  const int m0 = ti.tm * BM, n0 = ti.tn * BN;
  const bool full = (m0 + BM <= p.M) && (n0 + BN <= p.N) && !ti.tri_diag;
  // C and Cin may be the same array, so the compiler has to keep every load of Cin in front of the store that follows it:
  // written element by element, beta != 0 costs 64 (32 paired) dependent round trips to L2 per thread -- 32 us per tile,
  // measured (profiles/r02_update_probe.txt: time = K x ideal + 0.9 ms per 4096-tile launch at every K).  Hence batches:
  // all reads of a batch are in flight before its first write.
  if (!A_KC && p.c_vec2 && full) {
    // row tiles 2j and 2j+1 hold rows 16j+2g and 16j+2g+1: one 16-byte store per pair
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = m0 + acc_row<false>(wm, 2 * j, g);
#pragma unroll
      for (int jh = 0; jh < 4; jh += 2) {
        double2 o[2][2];
        if (beta != 0.0) {
#pragma unroll
          for (int jj = 0; jj < 2; ++jj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
              o[jj][e] = *reinterpret_cast<const double2*>(p.Cin + r + (int64_t)(n0 + acc_col<B_KC>(wn, jh + jj, 2 * t + e)) * p.ldcin);
        }
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int c = n0 + acc_col<B_KC>(wn, jh + jj, 2 * t + e);
            double2 v = make_double2(alpha * acc[2 * j][jh + jj][e], alpha * acc[2 * j + 1][jh + jj][e]);
            if (beta != 0.0) { v.x += beta * o[jj][e].x; v.y += beta * o[jj][e].y; }
            *reinterpret_cast<double2*>(C + r + (int64_t)c * p.ldc) = v;
          }
      }
    }
    return;
  }
  constexpr int RB = 2;                      // row tiles per batch: 32 reads in flight (the accumulators hold 128 of the 240 registers)
#pragma unroll
  for (int i0 = 0; i0 < 8; i0 += RB) {
    double o[RB][4][2];
    uint32_t live = 0;                       // bit (ii*8 + jj*2 + e): the element exists and is stored
#pragma unroll
    for (int ii = 0; ii < RB; ++ii) {
      const int r = m0 + acc_row<A_KC>(wm, i0 + ii, g);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int c = n0 + acc_col<B_KC>(wn, jj, 2 * t + e);
          bool ok = (r < p.M) && (c < p.N);
          if (ti.tri_diag && p.c_tri == CTRI_UPPER && r - m0 > c - n0) ok = false;     // (a triangular C crosses a tile only where
          if (ti.tri_diag && p.c_tri == CTRI_LOWER && r - m0 < c - n0) ok = false;     //  its row and column origins coincide)
          live |= (ok ? 1u : 0u) << (ii * 8 + jj * 2 + e);
          o[ii][jj][e] = (ok && beta != 0.0) ? p.Cin[r + (int64_t)c * p.ldcin] : 0.0;
        }
    }
#pragma unroll
    for (int ii = 0; ii < RB; ++ii) {
      const int r = m0 + acc_row<A_KC>(wm, i0 + ii, g);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int c = n0 + acc_col<B_KC>(wn, jj, 2 * t + e);
          if ((live >> (ii * 8 + jj * 2 + e)) & 1u) C[r + (int64_t)c * p.ldc] = fma(beta, o[ii][jj][e], alpha * acc[i0 + ii][jj][e]);
        }
    }
  }
}

// ------------------------------------------------------------------------------------------ kernel
// A_KC: op(A) is K-contiguous in global memory (transa = 'T'); B_KC: op(B) is K-contiguous (transb = 'N').
// mapA / mapB: 2-D tensor maps, dim0 = the contiguous index.  Boxes: K-contiguous {16 k, 128 mn};
// MN-contiguous {16 mn, 16 k} (issued 8 times per stage).
template <bool A_KC, bool B_KC>
__global__ void __launch_bounds__(T_THREADS, 1)
dmma_tma_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const TmaGemmParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* tiles = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + T_STAGES * T_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + T_STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < T_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], T_CONSUMER_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  int stage = 0;
  uint32_t phase = 0;

  if (warp >= T_CONSUMER_WARPS) {
    // ===================== producer: one lane feeds the ring, running ahead across tiles =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(T_REGS_PRODUCER));
    if (warp == T_CONSUMER_WARPS && lane == 0) {
      for (int unit = blockIdx.x; unit < p.units; unit += gridDim.x)
        for (int sub = 0; sub < p.subs; ++sub) {
          const TileInfo ti = decode_tile(p, unit, sub);
          if (tile_outside_tri(p, ti)) continue;
          const int m0 = ti.tm * BM, n0 = ti.tn * BN;
          for (int kt = 0; kt < ti.nk; ++kt) {
            mbar_wait(&empty_bar[stage], phase ^ 1u);
            mbar_expect_tx(&full_bar[stage], T_STAGE_BYTES);
            unsigned char* sA = tiles + stage * T_STAGE_BYTES;
            unsigned char* sB = sA + T_TILE_BYTES;
            const int k0 = ti.k_first + kt * ti.k_step;
            if (A_KC) tma_load_2d(sA, &mapA, k0, m0, &full_bar[stage]);
            else {
#pragma unroll
              for (int b = 0; b < 8; ++b) tma_load_2d(sA + b * 2048, &mapA, m0 + 16 * b, k0, &full_bar[stage]);
            }
            if (B_KC) tma_load_2d(sB, &mapB, k0, n0, &full_bar[stage]);
            else {
#pragma unroll
              for (int b = 0; b < 8; ++b) tma_load_2d(sB + b * 2048, &mapB, n0 + 16 * b, k0, &full_bar[stage]);
            }
            if (++stage == T_STAGES) { stage = 0; phase ^= 1u; }
          }
        }
    }
    return;
  }

  // ===================== consumers =====================
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(T_REGS_CONSUMER));
  const int wm = warp >> 2, wn = warp & 3, g = lane >> 2, t = lane & 3;
  const FragAddr fa = frag_addr<A_KC, B_KC>(wm, wn, g, t);
  const uint32_t tiles_u32 = smem_u32(tiles);
  for (int unit = blockIdx.x; unit < p.units; unit += gridDim.x)
    for (int sub = 0; sub < p.subs; ++sub) {
      const TileInfo ti = decode_tile(p, unit, sub);
      if (tile_outside_tri(p, ti)) continue;
      double acc[8][4][2];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) acc[i][jj][0] = acc[i][jj][1] = 0.0;
      const uint32_t tile_mask = ti.tri_diag ? ctri_mask(p.c_tri, wm, wn) : 0xFFu;
      // software pipeline in quarter stages (half of the k-columns x half of the rows): the next quarter's
      // fragments are in flight while the DMMAs of the current one issue (the 8 warps run in lockstep, so
      // un-overlapped LDS bursts would idle the pipe)
      FragA a0, a1;
      FragB b0, b1;
      if (ti.nk > 0) {
        mbar_wait(&full_bar[stage], phase);
        const uint32_t sb = tiles_u32 + stage * T_STAGE_BYTES;
        load_b<B_KC, 0>(sb, fa, b0);
        load_a<A_KC, 0, 0>(sb, fa, a0);
      }
      // k-stages [d0, d1) touch the diagonal block of the triangular operand (or the whole tile is a triangular
      // C tile): they run in their own loop so that the common loop carries no mask logic at all
      int d0 = ti.nk, d1 = ti.nk;
      if (tile_mask != 0xFFu) { d0 = 0; }
      else if (p.diag != DIAG_NONE && ti.nk > 0) {
        // stage kt covers k0 = k_first + kt * k_step; masked iff diag_origin <= k0 < diag_origin + BM
        const int lo = ti.diag_origin, hi = ti.diag_origin + BM - BK;   // first / last masked k0
        const int ka = ti.k_first, kb = ti.k_first + (ti.nk - 1) * ti.k_step;
        const int kmin = ti.k_step > 0 ? ka : kb, kmax = ti.k_step > 0 ? kb : ka;
        if (kmax >= lo && kmin <= hi) {
          const int first_k = lo > kmin ? lo : kmin, last_k = hi < kmax ? hi : kmax;   // masked k0 range (inclusive)
          if (ti.k_step > 0) { d0 = (first_k - ka) / BK; d1 = (last_k - ka) / BK + 1; }
          else { d0 = (ka - last_k) / BK; d1 = (ka - first_k) / BK + 1; }
        }
      }
      auto stages = [&](auto masked_tag, int kt0, int kt1) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        for (int kt = kt0; kt < kt1; ++kt) {
          uint32_t mask = tile_mask;
          if (MASKED && p.diag != DIAG_NONE) {
            const int k0 = ti.k_first + kt * ti.k_step;
            const int ds = (k0 - ti.diag_origin) / BK;   // 16-row step inside the diagonal block
            if (k0 >= ti.diag_origin && ds < BM / BK) mask &= diag_mask(p.diag, ds, wm, wn);
          }
          const uint32_t sb = tiles_u32 + stage * T_STAGE_BYTES;
          load_a<A_KC, 0, 1>(sb, fa, a1);
          mma_part<MASKED, 0>(acc, a0, b0, mask);
          load_b<B_KC, 1>(sb, fa, b1);
          load_a<A_KC, 1, 0>(sb, fa, a0);
          mma_part<MASKED, 1>(acc, a1, b0, mask);
          load_a<A_KC, 1, 1>(sb, fa, a1);
          mma_part<MASKED, 0>(acc, a0, b1, mask);
          const int cur = stage;
          if (++stage == T_STAGES) { stage = 0; phase ^= 1u; }
          if (kt + 1 < ti.nk) {
            mbar_wait(&full_bar[stage], phase);
            const uint32_t sn = tiles_u32 + stage * T_STAGE_BYTES;
            load_b<B_KC, 0>(sn, fa, b0);
            load_a<A_KC, 0, 0>(sn, fa, a0);
          }
          mma_part<MASKED, 1>(acc, a1, b1, mask);
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty_bar[cur]);
        }
      };
      stages(FalseTag{}, 0, d0);
      stages(TrueTag{}, d0, d1);
      stages(FalseTag{}, d1, ti.nk);
      store_tile<A_KC, B_KC>(p, ti, acc, wm, wn, g, t);
    }
}

}  // namespace capk
