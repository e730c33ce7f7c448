// C ABI (include/capital_b200.h) over the sm_100a kernels.  No CPU fallback anywhere in this file.
#include "../../include/capital_b200.h"
#include "aux_kernels.cuh"
#include "dmma_gemm.cuh"
#include "dmma_tma_gemm.cuh"
#include "handle.h"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

using namespace capk;

static thread_local std::string g_last_error;
const char* cap_last_error(void) { return g_last_error.c_str(); }
int cap_fail(int code, const char* what, const char* detail) {
  g_last_error = std::string(what) + (detail ? std::string(": ") + detail : std::string());
  return code;
}
#define CUDA_TRY(expr)                                                                   \
  do {                                                                                   \
    cudaError_t e_ = (expr);                                                             \
    if (e_ != cudaSuccess) return cap_fail(CAP_ERR_CUDA, #expr, cudaGetErrorString(e_)); \
  } while (0)
#define CAP_TRY(expr)            \
  do {                           \
    int rc_ = (expr);            \
    if (rc_ != CAP_OK) return rc_; \
  } while (0)
#define LAUNCH_CHECK(h)                                                                            \
  do {                                                                                             \
    (h)->launches++;                                                                               \
    cudaError_t e_ = cudaGetLastError();                                                           \
    if (e_ != cudaSuccess) return cap_fail(CAP_ERR_CUDA, "kernel launch", cudaGetErrorString(e_)); \
  } while (0)

static int noalloc_setup(cap_handle_s* h);   // defined with cap_cqr_gram_noalloc at the end of this file

// ------------------------------------------------------------------------------------------ context
int cap_create(cap_handle_t* out, int device) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return cap_fail(CAP_ERR_CUDA, "cap_create", "no CUDA device (there is no CPU fallback)");
  if (device < 0 || device >= count) return cap_fail(CAP_ERR_ARG, "cap_create", "bad device index");
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return cap_fail(CAP_ERR_UNSUPPORTED, "cap_create", "kernels are built for sm_100a only");
  cap_handle_s* h = new cap_handle_s();
  h->device = device;
  h->num_sms = prop.multiProcessorCount;
  if (const char* e = getenv("CAPITAL_B200_GRID_MULT")) h->grid_mult = std::max(1, atoi(e));   // probe switch: tile-engine grid = k x #SMs everywhere
  if (const char* e = getenv("CAPITAL_B200_STAIR")) h->lookahead_stair = atoi(e);
  if (const char* e = getenv("CAPITAL_B200_SUMMA_PIPE")) h->summa_pipeline = atoi(e);           // A/B switches of the pipelined SUMMA
  if (const char* e = getenv("CAPITAL_B200_SUMMA_RESERVE")) h->summa_reserve_sms = atoi(e);
  if (const char* e = getenv("CAPITAL_B200_SUMMA_HALVES")) h->summa_halves = atoi(e);     // A/B switch: 0 = trailing update block run by block run
  CUDA_TRY(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
  CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
  h->stream = h->own_stream;
  CUDA_TRY(cudaMalloc(&h->info_dev, sizeof(int)));
  CUDA_TRY(cudaMemset(h->info_dev, 0, sizeof(int)));
  // opt in to the large dynamic shared memory the tile engine and the base case need
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<false, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<false, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<true, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<true, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<false, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<false, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<true, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_gemm_kernel<true, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_trmm_right_inplace_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_trmm_right_inplace_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_trmm_right_inplace_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_trmm_right_inplace_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(potrf_trtri_base_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BASE_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_small_gemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_small_gemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_small_gemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_small_gemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_tma_gemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, T_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_tma_gemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, T_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_tma_gemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, T_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(dmma_tma_gemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, T_SMEM_BYTES));
  {  // cuTensorMapEncodeTiled through the runtime (no link-time dependency on libcuda)
    cudaDriverEntryPointQueryResult qres;
    void* fn = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
      cudaGetLastError();
      return cap_fail(CAP_ERR_CUDA, "cap_create", "cuTensorMapEncodeTiled is not available from this driver");
    }
    h->encode_tiled = fn;
  }
  CAP_TRY(noalloc_setup(h));     // address of the static scratch of cap_cqr_gram_noalloc on this device (no allocation)
  *out = h;
  return CAP_OK;
}

int cap_destroy(cap_handle_t h) {
  if (!h) return CAP_OK;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  for (auto& w : h->ws) if (w.ptr) cudaFree(w.ptr);
  if (h->info_dev) cudaFree(h->info_dev);
  if (h->summa_stream) {
    cudaStreamDestroy(h->summa_stream);
    for (cudaEvent_t e : {h->ev_sready[0], h->ev_sready[1], h->ev_sfree[0], h->ev_sfree[1], h->ev_shalf, h->ev_sred[0], h->ev_sred[1], h->ev_sstart}) cudaEventDestroy(e);
  }
  if (h->ev_a) cudaEventDestroy(h->ev_a);
  if (h->ev_b) cudaEventDestroy(h->ev_b);
  if (h->panel_stream) {
    cudaStreamDestroy(h->panel_stream);
    cudaStreamDestroy(h->gather_stream);
    for (cudaEvent_t e : {h->ev_panel[0], h->ev_panel[1], h->ev_u1, h->ev_trmm, h->ev_gath[0], h->ev_gath[1], h->ev_u2}) cudaEventDestroy(e);
  }
  if (h->xchg_stream) {
    cudaStreamDestroy(h->xchg_stream); cudaStreamDestroy(h->sig_stream); cudaStreamDestroy(h->row_stream);
    for (int i = 0; i < cap_handle_s::RING; ++i)
      for (cudaEvent_t e : {h->ev_leaf[i], h->ev_csolve[i], h->ev_solve[i], h->ev_xchg[i], h->ev_xpush[i], h->ev_rowready[i], h->ev_r1[i], h->ev_r2a[i], h->ev_r2[i]}) cudaEventDestroy(e);
    cudaFree(h->sig_scratch);
  }
  if (h->peer1d) { cudaDeviceSynchronize(); cap_peer_free(h->peer1d); h->peer1d = nullptr; }
  if (h->dup_comm) cap_comm_destroy(h->dup_comm);
  cudaStreamDestroy(h->own_stream);
  cudaStreamDestroy(h->copy_stream);
  delete h;
  return CAP_OK;
}

int cap_set_stream(cap_handle_t h, void* s) { h->stream = (cudaStream_t)s; return CAP_OK; }   // NULL = the legacy default stream
int cap_use_own_stream(cap_handle_t h) { h->stream = h->own_stream; return CAP_OK; }
void* cap_get_stream(cap_handle_t h) { return (void*)h->stream; }
int cap_synchronize(cap_handle_t h) { CUDA_TRY(cudaStreamSynchronize(h->stream)); return CAP_OK; }
long long cap_kernel_launches(cap_handle_t h) { return h->launches; }
int cap_get_info(cap_handle_t h, int* info) {
  CUDA_TRY(cudaMemcpyAsync(info, h->info_dev, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return CAP_OK;
}
int cap_reset_info(cap_handle_t h) { CUDA_TRY(cudaMemsetAsync(h->info_dev, 0, sizeof(int), h->stream)); return CAP_OK; }

// Library-owned scratch: slot-indexed, grown on demand, never shrunk (stream-ordered reuse on one stream).
int cap_workspace(cap_handle_t h, int slot, size_t bytes, double** out) {
  auto& w = h->ws[slot];
  if (w.bytes < bytes) {
    if (w.ptr) { CUDA_TRY(cudaDeviceSynchronize()); CUDA_TRY(cudaFree(w.ptr)); w.ptr = nullptr; w.bytes = 0; }   // every stream of the handle may be using the slot
    if (cudaMalloc(&w.ptr, bytes) != cudaSuccess) { cudaGetLastError(); return cap_fail(CAP_ERR_ALLOC, "workspace", "cudaMalloc failed"); }
    w.bytes = bytes;
  }
  *out = (double*)w.ptr;
  return CAP_OK;
}

int cap_release_workspace(cap_handle_t h) {
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaDeviceSynchronize());
  for (auto& w : h->ws) if (w.ptr) { CUDA_TRY(cudaFree(w.ptr)); w.ptr = nullptr; w.bytes = 0; }
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ memory
int cap_malloc(void** p, size_t bytes) {
  if (cudaMalloc(p, bytes) != cudaSuccess) { cudaGetLastError(); return cap_fail(CAP_ERR_ALLOC, "cap_malloc", "cudaMalloc failed"); }
  return CAP_OK;
}
int cap_free(void* p) { CUDA_TRY(cudaFree(p)); return CAP_OK; }
int cap_memset_async(cap_handle_t h, void* p, int v, size_t bytes) { CUDA_TRY(cudaMemsetAsync(p, v, bytes, h->stream)); return CAP_OK; }
int cap_h2d_async(cap_handle_t h, void* d, const void* s, size_t b) { CUDA_TRY(cudaMemcpyAsync(d, s, b, cudaMemcpyHostToDevice, h->stream)); return CAP_OK; }
int cap_d2h_async(cap_handle_t h, void* d, const void* s, size_t b) { CUDA_TRY(cudaMemcpyAsync(d, s, b, cudaMemcpyDeviceToHost, h->stream)); return CAP_OK; }
int cap_host_alloc(void** p, size_t bytes) {
  if (cudaHostAlloc(p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return cap_fail(CAP_ERR_ALLOC, "cap_host_alloc", "cudaHostAlloc failed"); }
  return CAP_OK;
}
int cap_host_free(void* p) { CUDA_TRY(cudaFreeHost(p)); return CAP_OK; }

// ------------------------------------------------------------------------------------------ tile engine dispatch
static inline bool aligned16(const void* p, int64_t ld) { return (((uintptr_t)p) & 15) == 0 && (ld & 1) == 0; }

template <bool TA, bool TB>
static void launch_gemm_t(cap_handle_t h, const GemmParams& p, int vec, dim3 grid) {
  if (vec == 2) dmma_gemm_kernel<TA, TB, 2><<<grid, NTHREADS, GEMM_SMEM_BYTES, h->stream>>>(p);
  else dmma_gemm_kernel<TA, TB, 1><<<grid, NTHREADS, GEMM_SMEM_BYTES, h->stream>>>(p);
}

// ---- TMA engine (dmma_tma_gemm.cuh) ----------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 2-D FP64 tensor map: dim0 = contiguous index (extent `inner`), dim1 strided by ld doubles (extent `outer`).
static int make_map(cap_handle_t h, CUtensorMap* map, const double* base, int64_t inner, int64_t outer, int64_t ld, bool k_contig) {
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
  cuuint32_t box[2] = {16u, k_contig ? 128u : 16u};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = ((encode_tiled_fn)h->encode_tiled)(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, estr,
                                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return cap_fail(CAP_ERR_CUDA, "cuTensorMapEncodeTiled", "tensor map rejected");
  return CAP_OK;
}

static bool tma_eligible(cap_handle_t h, const GemmParams& p) {
  return h->engine == 0 && p.K > 0 && aligned16(p.A, p.lda) && aligned16(p.B, p.ldb) &&
         p.lda < (int64_t(1) << 36) && p.ldb < (int64_t(1) << 36);
}

// split-K factor: outputs too small to fill the machine (tall-skinny Gram, small trailing updates) are cut
// into k-slabs.  Prefer a factor that makes the number of work units a multiple of the SM count (every
// persistent CTA then gets the same number of units, and with the triangular enumeration the same mix of
// diagonal and full tiles); otherwise one unit per SM.
static int pick_ksplit(cap_handle_t h, int ntiles, int64_t K, bool persistent) {
  if (ntiles * 2 > h->num_sms || K < 128) return 1;
  // long K: slabs of >= 1024 rows; short K (the small products inside cholinv / trtri, where one 128x128 tile
  // would otherwise sit on a single SM for K/16 x 2 us): slabs down to 4 pipeline stages
  const int64_t max_by_k = std::max<int64_t>(1, K >= 8192 ? K / 1024 : K / 64);
  if (persistent) {
    int a = ntiles, b = h->num_sms;
    while (b) { int tmp = a % b; a = b; b = tmp; }
    const int step = h->num_sms / a;                       // smallest s with ntiles*s % num_sms == 0
    if (step <= max_by_k && (int64_t)ntiles * step <= (int64_t)16 * h->num_sms) return step;
  }
  return (int)std::max<int64_t>(1, std::min<int64_t>(h->num_sms / ntiles, max_by_k));
}

template <bool A_KC, bool B_KC>
static void launch_tma_t(cap_handle_t h, const CUtensorMap& ma, const CUtensorMap& mb, const TmaGemmParams& tp, int grid) {
  dmma_tma_gemm_kernel<A_KC, B_KC><<<grid, T_THREADS, T_SMEM_BYTES, h->stream>>>(ma, mb, tp);
}

// C = alpha*op(A)*op(B) + beta*Cin with optional k-range skipping / triangular store / split-K.
// diag: DIAG_* of the triangular operand (its other triangle must hold zeros inside the diagonal blocks);
// rowpanel: one CTA owns a whole 128-row block of C and walks its column tiles in the order that makes
// B <- B*op(T) safe in place (also the balanced schedule for tall panels).
static int gemm_launch(cap_handle_t h, GemmParams p, bool ta, bool tb, int want_ksplit, int diag = DIAG_NONE, bool rowpanel = false) {
  if (p.M <= 0 || p.N <= 0) return CAP_OK;
  p.tiles_m = (p.M + BM - 1) / BM;
  p.tiles_n = (p.N + BN - 1) / BN;
  // small products (inner levels of cholinv / trtri): one warp per 32x32 block of C, operands straight from L2
  if (!rowpanel && h->engine == 0 && p.M <= SMALL_MAX_MN && p.N <= SMALL_MAX_MN && p.K <= 512 && (want_ksplit < 0 || want_ksplit == 1)) {
    const dim3 grid(((p.M + 31) / 32) * ((p.N + 31) / 32));
    p.ksplit = 1;
    if (ta) { if (tb) dmma_small_gemm_kernel<true, true><<<grid, SMALL_THREADS, SMALL_SMEM_BYTES, h->stream>>>(p); else dmma_small_gemm_kernel<true, false><<<grid, SMALL_THREADS, SMALL_SMEM_BYTES, h->stream>>>(p); }
    else    { if (tb) dmma_small_gemm_kernel<false, true><<<grid, SMALL_THREADS, SMALL_SMEM_BYTES, h->stream>>>(p); else dmma_small_gemm_kernel<false, false><<<grid, SMALL_THREADS, SMALL_SMEM_BYTES, h->stream>>>(p); }
    LAUNCH_CHECK(h);
    return CAP_OK;
  }
  const bool tri_sq = (p.c_tri != CTRI_FULL && p.tiles_m == p.tiles_n);
  int ntiles = tri_sq ? p.tiles_m * (p.tiles_m + 1) / 2 : p.tiles_m * p.tiles_n;
  const bool tma = tma_eligible(h, p);
  int ksplit = 1;
  if (want_ksplit < 0) ksplit = pick_ksplit(h, ntiles, p.K, tma); else ksplit = std::max(1, want_ksplit);
  double* user_c = p.C; const int64_t user_ldc = p.ldc; const double alpha = p.alpha, beta = p.beta;
  if (ksplit > 1) {
    int kper = (int)(((int64_t)p.K + ksplit - 1) / ksplit);
    kper = ((kper + BK - 1) / BK) * BK;
    ksplit = (p.K + kper - 1) / kper;
    if (ksplit > 1) {
      double* part;
      CAP_TRY(cap_workspace(h, h->ws_bank == 0 ? CAP_WS_GEMM_SPLITK : (h->ws_bank == 1 ? CAP_WS_GEMM_SPLITK2 : CAP_WS_GEMM_SPLITK3), (size_t)ksplit * p.M * p.N * sizeof(double), &part));
      p.k_per_split = kper; p.c_split_stride = (int64_t)p.M * p.N; p.C = part; p.ldc = p.M;
    }
  }
  p.ksplit = ksplit;
  if (tma) {
    CUtensorMap ma, mb;
    // op(A): 'T' -> stored K x M, k contiguous; 'N' -> stored M x K, m contiguous.  op(B): 'N' -> stored K x N, k contiguous.
    CAP_TRY(make_map(h, &ma, p.A, ta ? p.K : p.M, ta ? p.M : p.K, p.lda, ta));
    CAP_TRY(make_map(h, &mb, p.B, tb ? p.N : p.K, tb ? p.K : p.N, p.ldb, !tb));
    TmaGemmParams tp{};
    tp.C = p.C; tp.Cin = p.Cin; tp.ldc = p.ldc; tp.ldcin = p.ldcin; tp.M = p.M; tp.N = p.N; tp.K = p.K;
    tp.alpha = p.alpha; tp.beta = p.beta; tp.klo_mode = p.klo_mode; tp.khi_mode = p.khi_mode; tp.c_tri = p.c_tri;
    tp.diag = (ksplit > 1) ? DIAG_NONE : diag;
    tp.ksplit = ksplit; tp.k_per_split = p.k_per_split; tp.c_split_stride = p.c_split_stride;
    tp.tiles_m = p.tiles_m; tp.tiles_n = p.tiles_n;
    tp.subs = 1;
    if (rowpanel) {
      tp.sched = (p.khi_mode == KR_TILE_N) ? SCHED_ROWPANEL_DESC : SCHED_ROWPANEL_ASC;
      tp.units = p.tiles_m; tp.subs = p.tiles_n;
    } else if (tri_sq) { tp.sched = SCHED_TRI; tp.units = ntiles * ksplit; }
    else { tp.sched = (p.khi_mode == KR_TILE_M) ? SCHED_TM_DESC : SCHED_RASTER; tp.units = ntiles * ksplit; }
    tp.c_vec2 = (aligned16(tp.C, tp.ldc) && (beta == 0.0 || ksplit > 1 || aligned16(tp.Cin, tp.ldcin))) ? 1 : 0;
    const int grid = (int)std::min<int64_t>(tp.units, (int64_t)h->num_sms * std::max(1, h->grid_mult));
    const bool akc = ta, bkc = !tb;
    if (akc) { if (bkc) launch_tma_t<true, true>(h, ma, mb, tp, grid); else launch_tma_t<true, false>(h, ma, mb, tp, grid); }
    else     { if (bkc) launch_tma_t<false, true>(h, ma, mb, tp, grid); else launch_tma_t<false, false>(h, ma, mb, tp, grid); }
    LAUNCH_CHECK(h);
  } else if (rowpanel) {
    const int vec = (aligned16(p.A, p.lda) && aligned16(p.B, p.ldb)) ? 2 : 1;
    const int grid = p.tiles_m;
    if (ta) return cap_fail(CAP_ERR_ARG, "gemm_launch", "row-panel schedule needs op(A) = A");
    if (tb) { if (vec == 2) dmma_trmm_right_inplace_kernel<true, 2><<<grid, NTHREADS, GEMM_SMEM_BYTES, h->stream>>>(p);
              else dmma_trmm_right_inplace_kernel<true, 1><<<grid, NTHREADS, GEMM_SMEM_BYTES, h->stream>>>(p); }
    else    { if (vec == 2) dmma_trmm_right_inplace_kernel<false, 2><<<grid, NTHREADS, GEMM_SMEM_BYTES, h->stream>>>(p);
              else dmma_trmm_right_inplace_kernel<false, 1><<<grid, NTHREADS, GEMM_SMEM_BYTES, h->stream>>>(p); }
    LAUNCH_CHECK(h);
  } else {
    const int vec = (aligned16(p.A, p.lda) && aligned16(p.B, p.ldb)) ? 2 : 1;
    dim3 grid(ntiles, ksplit);
    if (ta) { if (tb) launch_gemm_t<true, true>(h, p, vec, grid); else launch_gemm_t<true, false>(h, p, vec, grid); }
    else    { if (tb) launch_gemm_t<false, true>(h, p, vec, grid); else launch_gemm_t<false, false>(h, p, vec, grid); }
    LAUNCH_CHECK(h);
  }
  if (ksplit > 1) {
    dim3 rg((p.M + 127) / 128, std::min(p.N, 65535));
    splitk_reduce_kernel<<<rg, 128, 0, h->stream>>>(p.C, p.c_split_stride, p.M, ksplit, user_c, user_ldc, p.M, p.N, alpha, beta, p.c_tri);
    LAUNCH_CHECK(h);
  }
  return CAP_OK;
}

// One launch for the whole trailing update of a rank in the 1-D look-ahead Cholesky (SCHED_STAIR, dmma_tma_gemm.cuh):
// C[:, local blocks r_first ..] (rows from global block row0) += alpha * A^T B on and above the global diagonal.
// A: K x M (the ordered row panel from block row0 on), B: K x N (this rank's solved blocks), both k-contiguous.
// Returns CAP_ERR_UNSUPPORTED when the operands do not meet TMA's alignment rules (the caller then goes block by block).
static int stair_launch(cap_handle_t h, const double* A, int64_t lda, const double* B, int64_t ldb, double* C, int64_t ldc,
                        int64_t M, int64_t N, int K, double alpha, int nb, int P, int me, int r_first, int row0) {
  if (M <= 0 || N <= 0) return CAP_OK;
  GemmParams g{}; g.A = A; g.B = B; g.lda = lda; g.ldb = ldb; g.K = K;
  if (!tma_eligible(h, g)) return CAP_ERR_UNSUPPORTED;
  TmaGemmParams tp{};
  tp.C = C; tp.Cin = C; tp.ldc = ldc; tp.ldcin = ldc; tp.M = (int)M; tp.N = (int)N; tp.K = K; tp.alpha = alpha; tp.beta = 1.0;
  tp.c_tri = CTRI_UPPER; tp.diag = DIAG_NONE; tp.ksplit = 1; tp.k_per_split = K;
  tp.tiles_m = (int)((M + BM - 1) / BM); tp.tiles_n = (int)((N + BN - 1) / BN);
  tp.sched = SCHED_STAIR; tp.subs = 1;
  tp.stair_T = nb / BM; tp.stair_P = P; tp.stair_me = me; tp.stair_r_first = r_first; tp.stair_row0 = row0;
  int64_t units = 0;
  const int T = tp.stair_T;
  for (int bi = 0; bi * T < tp.tiles_n; ++bi) {
    const int rb = r_first + bi;
    const int gblk = rb * P + ((rb & 1) ? P - 1 - me : me);
    const int Tb = std::min(T, tp.tiles_n - bi * T);
    const int base = (gblk - row0) * T + 1;
    if (base < 1) return cap_fail(CAP_ERR_ARG, "stair_launch", "block column left of the first row block");
    units += (int64_t)Tb * base + (int64_t)Tb * (Tb - 1) / 2;
  }
  if (units > INT32_MAX) return cap_fail(CAP_ERR_ARG, "stair_launch", "too many tiles");
  tp.units = (int)units;
  tp.c_vec2 = aligned16(C, ldc) ? 1 : 0;
  CUtensorMap ma, mb;
  CAP_TRY(make_map(h, &ma, A, K, M, lda, true));
  CAP_TRY(make_map(h, &mb, B, K, N, ldb, true));
  const int grid = (int)std::min<int64_t>(tp.units, (int64_t)h->num_sms * std::max(1, h->grid_mult));
  launch_tma_t<true, true>(h, ma, mb, tp, grid);
  LAUNCH_CHECK(h);
  return CAP_OK;
}

// Host-side view of the staircase enumeration (the very function the kernel decodes its work with), for CPU tests.
// Returns the number of work units of the launch geometry; unit >= 0 also decodes that unit.
int64_t cap_debug_stair_decode(int T, int P, int me, int r_first, int row0, int tiles_n, int64_t unit, int* tm, int* tn, int* tri) {
  int64_t units = 0;
  for (int bi = 0; bi * T < tiles_n; ++bi) {
    const int rb = r_first + bi;
    const int gblk = rb * P + ((rb & 1) ? P - 1 - me : me);
    const int Tb = std::min(T, tiles_n - bi * T);
    const int base = (gblk - row0) * T + 1;
    if (base < 1) return -1;
    units += (int64_t)Tb * base + (int64_t)Tb * (Tb - 1) / 2;
  }
  if (unit >= 0 && unit < units) {
    int a = 0, b = 0; bool t = false;
    capk::stair_decode(T, P, me, r_first, row0, tiles_n, (int)unit, a, b, t);
    if (tm) *tm = a; if (tn) *tn = b; if (tri) *tri = t ? 1 : 0;
  }
  return units;
}

int cap_set_host_chunk_bytes(cap_handle_t h, size_t bytes) {
  if (bytes < 4096) return cap_fail(CAP_ERR_ARG, "cap_set_host_chunk_bytes", "chunk too small");
  h->host_chunk_bytes = bytes;
  return CAP_OK;
}

int cap_set_cacqr_complete_inv(cap_handle_t h, int complete_inv) {
  h->cacqr_complete_inv = complete_inv ? 1 : 0;
  return CAP_OK;
}
int cap_set_peer_exchange(cap_handle_t h, int mode) {
  if (mode < -1 || mode > 1) return cap_fail(CAP_ERR_ARG, "cap_set_peer_exchange", "-1 = peer copy engines when CUDA IPC works, 0 = NCCL only, 1 = peer copy engines required");
  h->peer_exchange = mode;
  return CAP_OK;
}
int cap_set_update_grid(cap_handle_t h, int ctas_per_sm) {
  if (ctas_per_sm < 0 || ctas_per_sm > 64) return cap_fail(CAP_ERR_ARG, "cap_set_update_grid", "0 = persistent grid (one CTA per SM, SM-capped beside a chain), k > 0 = k x #SMs CTAs");
  h->lookahead_grid_mult = ctas_per_sm;
  return CAP_OK;
}

int cap_set_engine(cap_handle_t h, int engine) {
  if (engine != 0 && engine != 1) return cap_fail(CAP_ERR_ARG, "cap_set_engine", "0 = auto, 1 = cp.async engine only");
  h->engine = engine;
  return CAP_OK;
}

static inline bool is_t(char c) { return c == 'T' || c == 't' || c == 'C' || c == 'c'; }
static inline bool is_u(char c) { return c == 'U' || c == 'u'; }
static inline bool is_l(char c) { return c == 'L' || c == 'l'; }

int cap_dgemm(cap_handle_t h, char transa, char transb, int64_t m, int64_t n, int64_t k, double alpha,
              const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc) {
  if (m < 0 || n < 0 || k < 0 || m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_dgemm", "bad dimension");
  GemmParams p{};
  p.A = A; p.B = B; p.C = C; p.Cin = C; p.lda = lda; p.ldb = ldb; p.ldc = ldc; p.ldcin = ldc;
  p.M = (int)m; p.N = (int)n; p.K = (int)k; p.alpha = alpha; p.beta = beta;
  const int ntiles = (int)(((m + BM - 1) / BM) * ((n + BN - 1) / BN));
  (void)ntiles;
  return gemm_launch(h, p, is_t(transa), is_t(transb), -1);
}

int cap_dsyrk(cap_handle_t h, char uplo, char trans, int64_t n, int64_t k, double alpha,
              const double* A, int64_t lda, double beta, double* C, int64_t ldc) {
  if (n < 0 || k < 0 || n > INT32_MAX || k > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_dsyrk", "bad dimension");
  GemmParams p{};
  p.A = A; p.B = A; p.C = C; p.Cin = C; p.lda = lda; p.ldb = lda; p.ldc = ldc; p.ldcin = ldc;
  p.M = (int)n; p.N = (int)n; p.K = (int)k; p.alpha = alpha; p.beta = beta;
  p.c_tri = is_u(uplo) ? CTRI_UPPER : CTRI_LOWER;
  const int t = (int)((n + BM - 1) / BM);
  const bool tr = is_t(trans);   // 'T': C = A^T A (A is k x n);  'N': C = A A^T (A is n x k)
  (void)t;
  return gemm_launch(h, p, tr, !tr, -1);
}

__global__ void set_unit_diag_kernel(double* T, int64_t ld, int n, double v = 1.0) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) T[i + (int64_t)i * ld] = v;
}
int cap_fill_diagonal(cap_handle_t h, double* dst, int64_t ld, int64_t n, double value) {
  if (n <= 0) return CAP_OK;
  if (n > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_fill_diagonal", "bad dimension");
  set_unit_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(dst, ld, (int)n, value);
  LAUNCH_CHECK(h);
  return CAP_OK;
}

// C = alpha*op(T)*B (left) or alpha*B*op(T) (right).  t_clean: the other triangle of T is already zero.
static int trmm_impl(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
                     const double* T, int64_t ldt, const double* B, int64_t ldb, double* C, int64_t ldc, bool t_clean, bool inplace_right,
                     double beta = 0.0) {
  if (m < 0 || n < 0 || m > INT32_MAX || n > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_dtrmm", "bad dimension");
  if (m == 0 || n == 0) return CAP_OK;
  const bool left = is_l(side), upper = is_u(uplo), tr = is_t(trans), unit = is_u(diag);
  const int64_t nt = left ? m : n;
  if (!t_clean || unit) {  // masked copy of the referenced triangle (serialize-style), other half zero
    double* tc;
    CAP_TRY(cap_workspace(h, CAP_WS_TRI, (size_t)nt * nt * sizeof(double), &tc));
    launch_copy2d(h->stream, tc, nt, T, ldt, nt, nt, upper ? 3 : 4);
    LAUNCH_CHECK(h);
    if (unit) { set_unit_diag_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, h->stream>>>(tc, nt, (int)nt); LAUNCH_CHECK(h); }
    T = tc; ldt = nt;
  }
  GemmParams p{};
  p.C = C; p.Cin = C; p.ldc = ldc; p.ldcin = ldc; p.M = (int)m; p.N = (int)n; p.alpha = alpha; p.beta = beta;
  const bool lo_bound = (upper != tr);  // left: Upper/NoTrans or Lower/Trans -> k >= i ; right: those -> k <= j
  if (left) {
    p.A = T; p.lda = ldt; p.B = B; p.ldb = ldb; p.K = (int)m;
    if (lo_bound) p.klo_mode = KR_TILE_M; else p.khi_mode = KR_TILE_M;
    return gemm_launch(h, p, tr, false, -1, lo_bound ? DIAG_A_LE : DIAG_A_GE, false);
  }
  p.A = B; p.lda = ldb; p.B = T; p.ldb = ldt; p.K = (int)n;
  if (lo_bound) p.khi_mode = KR_TILE_N; else p.klo_mode = KR_TILE_N;
  // tall panels: one CTA per 128-row block walking all column tiles (balanced, and the only safe order in place)
  const bool rowpanel = inplace_right || (m + BM - 1) / BM >= 2 * (int64_t)h->num_sms;
  return gemm_launch(h, p, false, tr, rowpanel ? 1 : -1, lo_bound ? DIAG_B_LE : DIAG_B_GE, rowpanel);
}

int cap_dtrmm_oop(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
                  const double* T, int64_t ldt, const double* B, int64_t ldb, double* C, int64_t ldc) {
  return trmm_impl(h, side, uplo, trans, diag, m, n, alpha, T, ldt, B, ldb, C, ldc, false, false);
}

int cap_dtrmm(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
              const double* T, int64_t ldt, double* B, int64_t ldb) {
  if (!is_l(side)) return trmm_impl(h, side, uplo, trans, diag, m, n, alpha, T, ldt, B, ldb, B, ldb, false, true);
  if (m <= 0 || n <= 0) return CAP_OK;
  double* copy;  // left side: every output row block needs all of B's rows -> go through a copy of B
  CAP_TRY(cap_workspace(h, CAP_WS_PANEL, (size_t)m * n * sizeof(double), &copy));
  launch_copy2d(h->stream, copy, m, B, ldb, m, n, 0);
  LAUNCH_CHECK(h);
  return trmm_impl(h, side, uplo, trans, diag, m, n, alpha, T, ldt, copy, m, B, ldb, false, false);
}

int cap_cqr_gram(cap_handle_t h, int64_t m, int64_t n, const double* Q, int64_t ldq, double* G, int64_t ldg) {
  return cap_dsyrk(h, 'U', 'T', n, m, 1.0, Q, ldq, 0.0, G, ldg);
}

int cap_cqr_apply(cap_handle_t h, int64_t m, int64_t n, const double* Rinv, int64_t ldr,
                  const double* Qin, int64_t ldqin, double* Qout, int64_t ldqout) {
  const bool inplace = (Qin == Qout);
  if (inplace && ldqin != ldqout) return cap_fail(CAP_ERR_ARG, "cap_cqr_apply", "in-place call needs equal leading dimensions");
  return trmm_impl(h, 'R', 'U', 'N', 'N', m, n, 1.0, Rinv, ldr, Qin, ldqin, Qout, ldqout, /*t_clean=*/true, inplace);
}

// ------------------------------------------------------------------------------------------ LAPACK engine
static int memset2d(cap_handle_t h, double* p, int64_t ld, int64_t m, int64_t n) {
  CUDA_TRY(cudaMemset2DAsync(p, (size_t)ld * 8, 0, (size_t)m * 8, (size_t)n, h->stream));
  return CAP_OK;
}
static int copy2d(cap_handle_t h, double* dst, int64_t ldd, const double* src, int64_t lds, int64_t m, int64_t n, int mode) {
  if (m <= 0 || n <= 0) return CAP_OK;
  if (mode < 0 || mode > 4) return cap_fail(CAP_ERR_ARG, "cap_copy2d", "mode must be 0..4");
  launch_copy2d(h->stream, dst, ldd, src, lds, m, n, mode);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
static int base_case(cap_handle_t h, double* A, int64_t lda, double* X, int64_t ldx, int n, int info_base, int write_r, int skip_potrf, int lower) {
  potrf_trtri_base_kernel<<<1, BASE_THREADS, BASE_SMEM_BYTES, h->stream>>>(A, lda, X, ldx, n, h->info_dev, info_base, write_r, 1, skip_potrf, lower);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
// Every BASE_NB diagonal block of the n x n triangular T inverted in ONE launch (they are independent; inside the
// recursion they would be a serial chain of one-CTA kernels).  X: n x n with ldx (xblock == 0), or a row of compact
// xblock x xblock blocks with leading dimension xblock (block j = inverse of T's diagonal block j).
static int base_case_batched_inverse(cap_handle_t h, const double* T, int64_t ldt, double* X, int64_t ldx, int n, int lower, int xblock) {
  const int leaves = (n + BASE_NB - 1) / BASE_NB;
  potrf_trtri_base_kernel<<<leaves, BASE_THREADS, BASE_SMEM_BYTES, h->stream>>>(
      const_cast<double*>(T), ldt, X, ldx, n, h->info_dev, 0, 0, 1, 1, lower,
      xblock ? xblock / BASE_NB : 1 << 30, xblock ? (int64_t)xblock * xblock : 0);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
static inline int split_point(int len) { return ((len / 2 + BASE_NB - 1) / BASE_NB) * BASE_NB; }

// Recursive Cholesky + inverse on one device: the schedule of cholesky::cholinv::invoke
// (reference src/alg/cholesky/cholinv/cholinv.hpp:87-165) with a true triangular-store syrk.
// A(off.., off..) upper -> R in place; X = Rinv (zero-initialised by the caller).  W: 2 x (len/2)^2 scratch.
static int cholinv_rec(cap_handle_t h, double* A, int64_t lda, double* X, int64_t ldx, int off, int len, bool top, int complete_inv, double* W,
                       int info_off = 0) {
  double* A11 = A + off + (int64_t)off * lda;
  double* X11 = X + off + (int64_t)off * ldx;
  if (len <= BASE_NB) return base_case(h, A11, lda, X11, ldx, len, info_off + off, 1, 0, 0);
  const int n1 = split_point(len), n2 = len - n1;
  double* A12 = A11 + (int64_t)n1 * lda;
  double* A22 = A12 + n1;
  double* X12 = X11 + (int64_t)n1 * ldx;
  double* X22 = X12 + n1;
  CAP_TRY(cholinv_rec(h, A, lda, X, ldx, off, n1, false, complete_inv, W, info_off));
  // "trsm" (cholinv.hpp:116-121): R12 = R11^{-T} A12 through the explicit inverse
  CAP_TRY(trmm_impl(h, 'L', 'U', 'T', 'N', n1, n2, 1.0, X11, ldx, A12, lda, W, n1, true, false));
  CAP_TRY(copy2d(h, A12, lda, W, n1, n1, n2, 0));
  // trailing update (cholinv.hpp:131-134): A22 -= R12^T R12, upper triangle only
  CAP_TRY(cap_dsyrk(h, 'U', 'T', n2, n1, -1.0, W, n1, 1.0, A22, lda));
  CAP_TRY(cholinv_rec(h, A, lda, X, ldx, off + n1, n2, false, complete_inv, W, info_off));
  if (!(top && !complete_inv)) {
    // inverse completion (cholinv.hpp:147-156): Rinv12 = -(Rinv11 R12) Rinv22
    double* W2 = W + (int64_t)n1 * n2;
    CAP_TRY(trmm_impl(h, 'L', 'U', 'N', 'N', n1, n2, 1.0, X11, ldx, A12, lda, W2, n1, true, false));
    CAP_TRY(trmm_impl(h, 'R', 'U', 'N', 'N', n1, n2, -1.0, X22, ldx, W2, n1, X12, ldx, true, false));
  }
  return CAP_OK;
}

static size_t cholinv_scratch_doubles(int n) {
  if (n <= BASE_NB) return 2;
  const int n1 = split_point(n);
  return (size_t)2 * n1 * (n - n1) + 2;
}

int cap_dcholinv(cap_handle_t h, int n, double* A, int64_t lda, double* Rinv, int64_t ldri, int complete_inv) {
  if (n < 0) return cap_fail(CAP_ERR_ARG, "cap_dcholinv", "bad dimension");
  if (n == 0) return CAP_OK;
  double* W;
  CAP_TRY(cap_workspace(h, CAP_WS_PANEL, cholinv_scratch_doubles(n) * sizeof(double), &W));
  CAP_TRY(memset2d(h, Rinv, ldri, n, n));
  return cholinv_rec(h, A, lda, Rinv, ldri, 0, n, true, complete_inv, W);
}

// X = T^{-1} by recursive block inversion; T upper (lower: mirrored).  X zero-initialised by the caller.
// leaves_done: the BASE_NB diagonal blocks were already inverted (base_case_batched_inverse).
static int trtri_rec(cap_handle_t h, const double* T, int64_t ldt, double* X, int64_t ldx, int off, int len, bool lower, double* W,
                     bool leaves_done = false) {
  const double* T11 = T + off + (int64_t)off * ldt;
  double* X11 = X + off + (int64_t)off * ldx;
  if (len <= BASE_NB) return leaves_done ? CAP_OK : base_case(h, const_cast<double*>(T11), ldt, X11, ldx, len, off, 0, 1, lower ? 1 : 0);
  const int n1 = split_point(len), n2 = len - n1;
  CAP_TRY(trtri_rec(h, T, ldt, X, ldx, off, n1, lower, W, leaves_done));
  CAP_TRY(trtri_rec(h, T, ldt, X, ldx, off + n1, n2, lower, W, leaves_done));
  double* X22 = X11 + n1 + (int64_t)n1 * ldx;
  if (!lower) {  // X12 = -(X11 T12) X22
    const double* T12 = T11 + (int64_t)n1 * ldt;
    CAP_TRY(trmm_impl(h, 'L', 'U', 'N', 'N', n1, n2, 1.0, X11, ldx, T12, ldt, W, n1, true, false));
    CAP_TRY(trmm_impl(h, 'R', 'U', 'N', 'N', n1, n2, -1.0, X22, ldx, W, n1, X11 + (int64_t)n1 * ldx, ldx, true, false));
  } else {       // X21 = -(X22 T21) X11
    const double* T21 = T11 + n1;
    CAP_TRY(trmm_impl(h, 'L', 'L', 'N', 'N', n2, n1, 1.0, X22, ldx, T21, ldt, W, n2, true, false));
    CAP_TRY(trmm_impl(h, 'R', 'L', 'N', 'N', n2, n1, -1.0, X11, ldx, W, n2, X11 + n1, ldx, true, false));
  }
  return CAP_OK;
}

int cap_dtrtri(cap_handle_t h, char uplo, char diag, int n, double* A, int lda) {
  if (is_u(diag)) return cap_fail(CAP_ERR_UNSUPPORTED, "cap_dtrtri", "unit-diagonal inverse is not on the path");
  if (n <= 0) return n == 0 ? CAP_OK : cap_fail(CAP_ERR_ARG, "cap_dtrtri", "bad dimension");
  const bool lower = is_l(uplo);
  double *X, *W;
  CAP_TRY(cap_workspace(h, CAP_WS_FACTOR, (size_t)n * n * sizeof(double), &X));
  CAP_TRY(cap_workspace(h, CAP_WS_PANEL, cholinv_scratch_doubles(n) * sizeof(double), &W));
  CAP_TRY(memset2d(h, X, n, n, n));
  // the n/128 diagonal leaves are independent: one launch instead of a serial chain of one-CTA kernels
  CAP_TRY(base_case_batched_inverse(h, A, lda, X, n, n, lower ? 1 : 0, 0));
  CAP_TRY(trtri_rec(h, A, lda, X, n, 0, n, lower, W, true));
  return copy2d(h, A, lda, X, n, n, n, lower ? 2 : 1);   // only the referenced triangle is overwritten (LAPACK semantics)
}

// ------------------------------------------------------------------------------------------ triangular solve
// op(T) X = alpha B (left) or X op(T) = alpha B (right), in place on B.  Recursion on the triangular operand: solve with
// one diagonal half, update the other half of B with the off-diagonal block on the tile engine (the flops: a plain
// gemm), solve with the second half; a diagonal block of at most `nb` is applied through its explicit inverse
// (a triangular multiply on the same engine), so the whole solve is tensor-pipe work.  Not a reference engine call
// (src/blas/interface.h has no _trsm; the reference forms explicit inverses instead, cholinv.hpp:113-124) -- it is what
// makes cap_dpotrf an N^3/3 schedule.
struct TrsmCtx {
  cap_handle_t h;
  bool left, upper, trans;
  const double* T; int64_t ldt;   // the whole triangular operand; blocks are addressed by their diagonal offset
  const double* Xd; int nb;       // compact inverses of its nb x nb diagonal blocks: block j at Xd + j*nb*nb, ld nb
  double* W;                      // nb x (other dimension of B): staging of the left-side leaves
};
static inline int block_split(int len, int nb) { return ((len / 2 + nb - 1) / nb) * nb; }

// B: rows [off, off+len) of the right-hand side (left; len x other) or its columns [off, off+len) (right; other x len)
static int trsm_rec(const TrsmCtx& c, int off, int len, double* B, int64_t ldb, int64_t other) {
  cap_handle_t h = c.h;
  const char uplo = c.upper ? 'U' : 'L', tr = c.trans ? 'T' : 'N';
  if (len <= c.nb) {
    const double* Xi = c.Xd + (int64_t)(off / c.nb) * c.nb * c.nb;
    if (c.left) {
      CAP_TRY(copy2d(h, c.W, len, B, ldb, len, other, 0));
      return trmm_impl(h, 'L', uplo, tr, 'N', len, other, 1.0, Xi, c.nb, c.W, len, B, ldb, true, false);
    }
    return trmm_impl(h, 'R', uplo, tr, 'N', other, len, 1.0, Xi, c.nb, B, ldb, B, ldb, true, true);
  }
  const int n1 = block_split(len, c.nb), n2 = len - n1;
  const double* T11 = c.T + off + (int64_t)off * c.ldt;
  const double* Toff = c.upper ? T11 + (int64_t)n1 * c.ldt : T11 + n1;   // stored off-diagonal block: n1 x n2 (upper) or n2 x n1 (lower)
  const bool op_lower = (c.upper == c.trans);                            // op(T) is lower triangular
  if (c.left) {
    double *B1 = B, *B2 = B + n1;
    if (op_lower) {   // X1 = P11^-1 B1;  B2 -= P21 X1;  X2 = P22^-1 B2
      CAP_TRY(trsm_rec(c, off, n1, B1, ldb, other));
      CAP_TRY(cap_dgemm(h, tr, 'N', n2, other, n1, -1.0, Toff, c.ldt, B1, ldb, 1.0, B2, ldb));
      return trsm_rec(c, off + n1, n2, B2, ldb, other);
    }
    CAP_TRY(trsm_rec(c, off + n1, n2, B2, ldb, other));   // X2 = P22^-1 B2;  B1 -= P12 X2;  X1 = P11^-1 B1
    CAP_TRY(cap_dgemm(h, tr, 'N', n1, other, n2, -1.0, Toff, c.ldt, B2, ldb, 1.0, B1, ldb));
    return trsm_rec(c, off, n1, B1, ldb, other);
  }
  double *B1 = B, *B2 = B + (int64_t)n1 * ldb;
  if (!op_lower) {    // X1 = B1 P11^-1;  B2 -= X1 P12;  X2 = B2 P22^-1
    CAP_TRY(trsm_rec(c, off, n1, B1, ldb, other));
    CAP_TRY(cap_dgemm(h, 'N', tr, other, n2, n1, -1.0, B1, ldb, Toff, c.ldt, 1.0, B2, ldb));
    return trsm_rec(c, off + n1, n2, B2, ldb, other);
  }
  CAP_TRY(trsm_rec(c, off + n1, n2, B2, ldb, other));     // X2 = B2 P22^-1;  B1 -= X2 P21;  X1 = B1 P11^-1
  CAP_TRY(cap_dgemm(h, 'N', tr, other, n1, n2, -1.0, B2, ldb, Toff, c.ldt, 1.0, B1, ldb));
  return trsm_rec(c, off, n1, B1, ldb, other);
}

static int axpby2d(cap_handle_t h, double* y, int64_t ldy, const double* x, int64_t ldx, int64_t m, int64_t n, double alpha, double beta, int tri);

int cap_set_potrf_schedule(cap_handle_t h, int schedule) {
  if (schedule != 0 && schedule != 1) return cap_fail(CAP_ERR_ARG, "cap_set_potrf_schedule", "0 = right-looking with look-ahead, 1 = recursive");
  h->potrf_schedule = schedule;
  return CAP_OK;
}

int cap_set_block_sizes(cap_handle_t h, int potrf_nb, int trsm_nb) {
  if (potrf_nb < 0 || potrf_nb % BASE_NB || trsm_nb < BASE_NB || trsm_nb % BASE_NB)
    return cap_fail(CAP_ERR_ARG, "cap_set_block_sizes", "block sizes must be positive multiples of 128 (potrf_nb: 0 = chosen by size)");
  h->potrf_nb = potrf_nb; h->trsm_nb = trsm_nb;
  return CAP_OK;
}

int cap_dtrsm(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
              const double* T, int64_t ldt, double* B, int64_t ldb) {
  if (m < 0 || n < 0 || m > INT32_MAX || n > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_dtrsm", "bad dimension");
  if (is_u(diag)) return cap_fail(CAP_ERR_UNSUPPORTED, "cap_dtrsm", "unit-diagonal solve is not on the path");
  if (m == 0 || n == 0) return CAP_OK;
  if (alpha == 0.0) return memset2d(h, B, ldb, m, n);
  const bool left = is_l(side), lower = is_l(uplo);
  const int nt = (int)(left ? m : n), nb = h->trsm_nb;
  const int64_t other = left ? n : m;
  const int nblk = (nt + nb - 1) / nb;
  double *Xd, *W;
  const size_t inv_scratch = cholinv_scratch_doubles(std::min(nt, nb));
  CAP_TRY(cap_workspace(h, CAP_WS_FACTOR, (size_t)nblk * nb * nb * sizeof(double), &Xd));
  CAP_TRY(cap_workspace(h, CAP_WS_PANEL, std::max<size_t>(inv_scratch, left ? (size_t)std::min(nt, nb) * other : 0) * sizeof(double), &W));
  // inverses of the nb x nb diagonal blocks: all 128-leaves in one launch, then the block products of each block
  CUDA_TRY(cudaMemsetAsync(Xd, 0, (size_t)nblk * nb * nb * sizeof(double), h->stream));
  CAP_TRY(base_case_batched_inverse(h, T, ldt, Xd, nb, nt, lower ? 1 : 0, nb));
  for (int j = 0; j < nblk; ++j) {
    const int off = j * nb, len = std::min(nb, nt - off);
    if (len > BASE_NB) CAP_TRY(trtri_rec(h, T + off + (int64_t)off * ldt, ldt, Xd + (int64_t)j * nb * nb, nb, 0, len, lower, W, true));
  }
  if (alpha != 1.0) CAP_TRY(axpby2d(h, B, ldb, B, ldb, m, n, alpha, 0.0, CTRI_FULL));
  TrsmCtx c{h, left, !lower, is_t(trans), T, ldt, Xd, nb, W};
  return trsm_rec(c, 0, nt, B, ldb, other);
}

// lapack::engine::_potrf (src/lapack/interface.hpp:31-43), upper: A = R^T R in place, N^3/3 flops.
//   recursion:  potrf(A11);  R12 = R11^-T A12 (trsm_rec);  A22 -= R12^T R12 (triangular syrk);  potrf(A22)
// Diagonal blocks of at most potrf_nb are factored AND inverted by the cholinv recursion (the leaf kernel produces
// both anyway); those inverses are exactly the leaves trsm_rec needs, so no inverse larger than potrf_nb is ever formed
// (cap_dcholinv, the reference's schedule, forms them at every level: 5N^3/12 flops for the same R).
struct PotrfCtx { cap_handle_t h; double* A; int64_t lda; double* Xd; int nb; double* Wleaf; double* Wtrsm; };
static int potrf_rec(const PotrfCtx& c, int off, int len) {
  cap_handle_t h = c.h;
  double* A11 = c.A + off + (int64_t)off * c.lda;
  if (len <= c.nb)
    return cholinv_rec(h, A11, c.lda, c.Xd + (int64_t)(off / c.nb) * c.nb * c.nb, c.nb, 0, len, true, 1, c.Wleaf, off);
  const int n1 = block_split(len, c.nb), n2 = len - n1;
  double* A12 = A11 + (int64_t)n1 * c.lda;
  double* A22 = A12 + n1;
  CAP_TRY(potrf_rec(c, off, n1));
  TrsmCtx t{h, true, true, true, c.A, c.lda, c.Xd, c.nb, c.Wtrsm};
  CAP_TRY(trsm_rec(t, off, n1, A12, c.lda, n2));
  CAP_TRY(cap_dsyrk(h, 'U', 'T', n2, n1, -1.0, A12, c.lda, 1.0, A22, c.lda));
  return potrf_rec(c, off + n1, n2);
}

int cap_dpotrf(cap_handle_t h, char uplo, int n, double* A, int lda) {
  if (!is_u(uplo)) return cap_fail(CAP_ERR_UNSUPPORTED, "cap_dpotrf", "only uplo='U' (the reference asserts dir=='U', cholinv.hpp:9)");
  if (n <= 0) return n == 0 ? CAP_OK : cap_fail(CAP_ERR_ARG, "cap_dpotrf", "bad dimension");
  // block size: measured on B200 (tools/factor_probe.py, tools/potrf1d_check.py) 1024 is best at n = 8192, 2048 at 16384
  const int nb = h->potrf_nb > 0 ? h->potrf_nb : std::min(2048, std::max(512, (n / 8 + BASE_NB - 1) / BASE_NB * BASE_NB));
  if (n <= nb) {   // one block: the cholinv recursion without its top-level inverse completion
    double* X;
    CAP_TRY(cap_workspace(h, CAP_WS_FACTOR, (size_t)n * n * sizeof(double), &X));
    return cap_dcholinv(h, n, A, lda, X, n, 0);
  }
  if (h->potrf_schedule == 0) return cap_potrf_1d(h, nullptr, n, nb, A, lda);   // right-looking with look-ahead (potrf_1d.cuh)
  const int nblk = (n + nb - 1) / nb;
  const size_t leaf_scratch = cholinv_scratch_doubles(nb);
  double *Xd, *W;
  CAP_TRY(cap_workspace(h, CAP_WS_FACTOR, (size_t)nblk * nb * nb * sizeof(double), &Xd));
  CAP_TRY(cap_workspace(h, CAP_WS_PANEL, (leaf_scratch + (size_t)nb * n) * sizeof(double), &W));
  CUDA_TRY(cudaMemsetAsync(Xd, 0, (size_t)nblk * nb * nb * sizeof(double), h->stream));
  PotrfCtx c{h, A, lda, Xd, nb, W, W + leaf_scratch};
  return potrf_rec(c, 0, n);
}

// ------------------------------------------------------------------------------------------ serialize / util / generators
int cap_copy2d(cap_handle_t h, double* dst, int64_t ldd, const double* src, int64_t lds, int64_t m, int64_t n, int mode) {
  return copy2d(h, dst, ldd, src, lds, m, n, mode);
}
int cap_pack_upper(cap_handle_t h, double* packed, const double* full, int64_t ld, int n) {
  if (n <= 0) return CAP_OK;
  pack_upper_kernel<<<dim3((n + 255) / 256, std::min((n + PACK_COLS - 1) / PACK_COLS, 65535)), 256, 0, h->stream>>>(packed, full, ld, n);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
int cap_unpack_upper(cap_handle_t h, double* full, int64_t ld, const double* packed, int n, int zero_lower) {
  if (n <= 0) return CAP_OK;
  unpack_upper_kernel<<<dim3((n + 255) / 256, std::min((n + PACK_COLS - 1) / PACK_COLS, 65535)), 256, 0, h->stream>>>(full, ld, packed, n, zero_lower);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
int cap_axpby(cap_handle_t h, double* y, const double* x, int64_t count, double alpha, double beta) {
  if (count <= 0) return CAP_OK;
  const int blocks = (int)std::min<int64_t>((count + 255) / 256, (int64_t)h->num_sms * 16);
  axpby_kernel<<<blocks, 256, 0, h->stream>>>(y, x, count, alpha, beta);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
int cap_fill_random(cap_handle_t h, double* dst, int64_t ld, int64_t rows, int64_t cols, long long key) {
  const int64_t total = rows * cols;
  if (total <= 0) return CAP_OK;
  const int64_t threads = (total + 15) / 16;
  fill_random_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, h->stream>>>(dst, ld, rows, cols, key, 0);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
int cap_fill_symmetric(cap_handle_t h, double* dst, int64_t ld, int64_t rows, int64_t cols, int64_t px, int64_t py,
                       int64_t grid_x, int64_t grid_y, int64_t global_dim, int diagonally_dominant) {
  if (rows <= 0 || cols <= 0) return CAP_OK;
  dim3 g((unsigned)std::min<int64_t>((rows + 255) / 256, 65535), (unsigned)std::min<int64_t>(cols, 65535));
  fill_symmetric_kernel<<<g, 256, 0, h->stream>>>(dst, ld, rows, cols, px, py, grid_x, grid_y, global_dim, diagonally_dominant ? (double)global_dim : 0.0);
  LAUNCH_CHECK(h);
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ CholeskyQR2, 1D
// qr::cacqr::sweep_1d (reference src/alg/qr/cacqr/cacqr.hpp:7-29) with the Serialize policy's packed Gram
// exchange (policy.h:77-85): syrk -> pack -> allreduce -> unpack -> potrf+trtri -> Q = Qin * Rinv.
static int cqr_sweep(cap_handle_t h, cap_comm_t comm, int64_t m, int64_t n, const double* Qin, int64_t ldin, double* Qout, int64_t ldout,
                     double* G, double* Rinv, double* packed) {
  CAP_TRY(memset2d(h, G, n, n, n));
  CAP_TRY(cap_cqr_gram(h, m, n, Qin, ldin, G, n));
  if (comm) {
    CAP_TRY(cap_pack_upper(h, packed, G, n, (int)n));
    CAP_TRY(cap_allreduce_sum(h, comm, packed, n * (n + 1) / 2));
    CAP_TRY(cap_unpack_upper(h, G, n, packed, (int)n, 0));
  }
  CAP_TRY(cap_dcholinv(h, (int)n, G, n, Rinv, n, 1));
  return cap_cqr_apply(h, m, n, Rinv, n, Qin, ldin, Qout, ldout);
}

int cap_cacqr_factor_1d(cap_handle_t h, cap_comm_t comm, int64_t m_loc, int64_t n, const double* A, int64_t lda,
                        double* Q, int64_t ldq, double* R, int64_t ldr, int num_iter) {
  if (m_loc < 0 || n <= 0 || n > INT32_MAX || m_loc > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_1d", "bad dimension");
  if (num_iter < 1 || num_iter > 2) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_1d", "num_iter must be 1 or 2");
  double* ws;
  const size_t nn = (size_t)n * n;
  CAP_TRY(cap_workspace(h, CAP_WS_GRAM, (4 * nn) * sizeof(double), &ws));
  double *G1 = ws, *G2 = ws + nn, *Rinv = ws + 2 * nn, *packed = ws + 3 * nn;
  CAP_TRY(cqr_sweep(h, comm, m_loc, n, A, lda, Q, ldq, G1, Rinv, packed));          // A is never copied: sweep 1 reads A, writes Q
  if (num_iter == 1) return copy2d(h, R, ldr, G1, n, n, n, 3);
  CAP_TRY(cqr_sweep(h, comm, m_loc, n, Q, ldq, Q, ldq, G2, Rinv, packed));          // sweep 2 in place on Q
  // R = R2 * R1 (cacqr.hpp:184-187); both factors have zero strict lower parts here
  return trmm_impl(h, 'R', 'U', 'N', 'N', n, n, 1.0, G1, n, G2, n, R, ldr, true, false);
}

// End to end from HOST memory, pipelined over row chunks on two streams:
//   sweep 1: H2D of chunk i+1 (copy stream) overlaps the Gram of chunk i (G1 += A_i^T A_i);
//   sweep 2: the in-place apply of chunk i overlaps the D2H of chunk i-1.
// The step is PCIe-bound (2 x 8 m n bytes); what stays exposed is apply 1 + Gram 2 + the two small factorizations.
int cap_cacqr_factor_1d_host(cap_handle_t h, cap_comm_t comm, int64_t m_loc, int64_t n, const double* A_host, int64_t lda,
                             double* Q_host, int64_t ldq, double* R_host, int64_t ldr, int num_iter) {
  if (lda != m_loc || ldq != m_loc || ldr != n) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_1d_host", "host buffers must be dense (ld == rows)");
  if (m_loc < 0 || n <= 0 || n > INT32_MAX || m_loc > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_1d_host", "bad dimension");
  if (num_iter < 1 || num_iter > 2) return cap_fail(CAP_ERR_ARG, "cap_cacqr_factor_1d_host", "num_iter must be 1 or 2");
  const size_t bytes = (size_t)m_loc * n * sizeof(double);
  const size_t nn = (size_t)n * n;
  double *dA, *dQ, *dR, *ws;
  CAP_TRY(cap_workspace(h, CAP_WS_STAGE0, bytes, &dA));
  CAP_TRY(cap_workspace(h, CAP_WS_STAGE1, bytes, &dQ));
  CAP_TRY(cap_workspace(h, CAP_WS_FACTOR, nn * sizeof(double), &dR));
  CAP_TRY(cap_workspace(h, CAP_WS_GRAM, (4 * nn) * sizeof(double), &ws));
  double *G1 = ws, *G2 = ws + nn, *Rinv = ws + 2 * nn, *packed = ws + 3 * nn;
  if (!h->ev_a) { CUDA_TRY(cudaEventCreateWithFlags(&h->ev_a, cudaEventDisableTiming)); CUDA_TRY(cudaEventCreateWithFlags(&h->ev_b, cudaEventDisableTiming)); }
  // chunks of ~1 GiB, multiples of 128 rows (tile height), at most 32
  // (and at least 16 chunks when the block is large enough for chunks of >= 64 MiB: with few chunks per rank -- 4 at 8 GPUs in
  //  round 1 -- the un-overlapped first upload and last download are a visible part of the step)
  const int64_t chunk_bytes = std::min<int64_t>((int64_t)h->host_chunk_bytes, std::max<int64_t>((int64_t)64 << 20, (int64_t)bytes / 16));
  int64_t rows_per = ((std::max<int64_t>(1, chunk_bytes / (n * 8)) + 127) / 128) * 128;
  if ((m_loc + rows_per - 1) / rows_per > 32) rows_per = (((m_loc + 31) / 32 + 127) / 128) * 128;
  const int nchunks = m_loc > 0 ? (int)((m_loc + rows_per - 1) / rows_per) : 0;
  cudaStream_t cs = h->copy_stream;
  CUDA_TRY(cudaEventRecord(h->ev_b, h->stream));            // staging buffers may still be in use by earlier work
  CUDA_TRY(cudaStreamWaitEvent(cs, h->ev_b, 0));
  CAP_TRY(memset2d(h, G1, n, n, n));
  for (int c = 0; c < nchunks; ++c) {
    const int64_t r0 = c * rows_per, rows = std::min<int64_t>(rows_per, m_loc - r0);
    CUDA_TRY(cudaMemcpy2DAsync(dA + r0, (size_t)m_loc * 8, A_host + r0, (size_t)lda * 8, (size_t)rows * 8, (size_t)n, cudaMemcpyHostToDevice, cs));
    CUDA_TRY(cudaEventRecord(h->ev_a, cs));
    CUDA_TRY(cudaStreamWaitEvent(h->stream, h->ev_a, 0));
    CAP_TRY(cap_dsyrk(h, 'U', 'T', n, rows, 1.0, dA + r0, m_loc, c == 0 ? 0.0 : 1.0, G1, n));
  }
  auto finish_gram = [&](double* G) -> int {
    if (comm) {
      CAP_TRY(cap_pack_upper(h, packed, G, n, (int)n));
      CAP_TRY(cap_allreduce_sum(h, comm, packed, n * (n + 1) / 2));
      CAP_TRY(cap_unpack_upper(h, G, n, packed, (int)n, 0));
    }
    return cap_dcholinv(h, (int)n, G, n, Rinv, n, 1);
  };
  CAP_TRY(finish_gram(G1));
  if (num_iter == 1) {
    CAP_TRY(copy2d(h, dR, n, G1, n, n, n, 3));
  } else {
    CAP_TRY(cap_cqr_apply(h, m_loc, n, Rinv, n, dA, m_loc, dQ, m_loc));
    CAP_TRY(memset2d(h, G2, n, n, n));
    CAP_TRY(cap_cqr_gram(h, m_loc, n, dQ, m_loc, G2, n));
    CAP_TRY(finish_gram(G2));
    CAP_TRY(trmm_impl(h, 'R', 'U', 'N', 'N', n, n, 1.0, G1, n, G2, n, dR, n, true, false));
  }
  const double* src = (num_iter == 1) ? dA : dQ;
  for (int c = 0; c < nchunks; ++c) {
    const int64_t r0 = c * rows_per, rows = std::min<int64_t>(rows_per, m_loc - r0);
    CAP_TRY(cap_cqr_apply(h, rows, n, Rinv, n, src + r0, m_loc, dQ + r0, m_loc));
    CUDA_TRY(cudaEventRecord(h->ev_b, h->stream));
    CUDA_TRY(cudaStreamWaitEvent(cs, h->ev_b, 0));
    CUDA_TRY(cudaMemcpy2DAsync(Q_host + r0, (size_t)ldq * 8, dQ + r0, (size_t)m_loc * 8, (size_t)rows * 8, (size_t)n, cudaMemcpyDeviceToHost, cs));
  }
  CUDA_TRY(cudaMemcpyAsync(R_host, dR, nn * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  CUDA_TRY(cudaStreamSynchronize(cs));
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ process-grid algorithms
// ------------------------------------------------------------------------------------------ validators
// util::residual_local (src/util/util.hpp:27-53) and the metrics built on it (test/qr/validate.hpp:7-62,
// test/cholesky/validate.hpp:7-30), computed on the device with the library's own products.
static int resid_ws(cap_handle_t h, double** acc, double** partial) {
  double* ws;
  CAP_TRY(cap_workspace(h, CAP_WS_VALID, (size_t)(2 + 2 * RESID_BLOCKS) * sizeof(double), &ws));
  *acc = ws; *partial = ws + 2;
  return CAP_OK;
}
int cap_residual_begin(cap_handle_t h) {
  double *acc, *partial;
  CAP_TRY(resid_ws(h, &acc, &partial));
  CUDA_TRY(cudaMemsetAsync(acc, 0, 2 * sizeof(double), h->stream));
  return CAP_OK;
}
int cap_residual_accumulate(cap_handle_t h, int64_t m, int64_t n, const double* X, int64_t ldx, const double* Ref, int64_t ldref,
                            int d, int x, int y, int upper_only, double offdiag_weight) {
  if (m <= 0 || n <= 0) return CAP_OK;
  double *acc, *partial;
  CAP_TRY(resid_ws(h, &acc, &partial));
  residual_partial_kernel<<<RESID_BLOCKS, RESID_THREADS, 0, h->stream>>>(X, ldx, Ref, ldref, m, n, d, x, y, upper_only, offdiag_weight, partial);
  LAUNCH_CHECK(h);
  residual_final_kernel<<<1, 32, 0, h->stream>>>(partial, acc);
  LAUNCH_CHECK(h);
  return CAP_OK;
}
int cap_residual_finish(cap_handle_t h, cap_comm_t comm, double* out) {
  double *acc, *partial, host[2];
  CAP_TRY(resid_ws(h, &acc, &partial));
  if (comm) CAP_TRY(cap_allreduce_sum(h, comm, acc, 2));
  CUDA_TRY(cudaMemcpyAsync(host, acc, sizeof(host), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  *out = host[1] > 0.0 ? std::sqrt(host[0]) / std::sqrt(host[1]) : std::sqrt(host[0]);
  return CAP_OK;
}

// qr::validate::orthogonality / residual for the 1D algorithm: orth = ||Q^T Q - I||_F / n, resid = ||Q R - A||_F / ||A||_F.
int cap_validate_cacqr_1d(cap_handle_t h, cap_comm_t comm, int64_t m_loc, int64_t n, const double* A, int64_t lda,
                          const double* Q, int64_t ldq, const double* R, int64_t ldr, double* orth, double* resid) {
  if (m_loc < 0 || n <= 0 || n > INT32_MAX || m_loc > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_validate_cacqr_1d", "bad dimension");
  double* G;
  CAP_TRY(cap_workspace(h, CAP_WS_GRAM, (size_t)4 * n * n * sizeof(double), &G));
  CAP_TRY(memset2d(h, G, n, n, n));
  if (m_loc > 0) CAP_TRY(cap_dsyrk(h, 'U', 'T', n, m_loc, 1.0, Q, ldq, 0.0, G, n));
  if (comm) CAP_TRY(cap_allreduce_sum(h, comm, G, n * n));
  CAP_TRY(cap_residual_begin(h));
  CAP_TRY(cap_residual_accumulate(h, n, n, G, n, nullptr, 0, 1, 0, 0, 1, 2.0));     // upper triangle formed: off-diagonal counts twice
  CAP_TRY(cap_residual_finish(h, nullptr, orth));
  // Q R - A over row chunks of at most ~1 GiB
  const int64_t rows_per = std::max<int64_t>(128, ((((int64_t)1 << 27) / n) / 128) * 128);
  double* T;
  CAP_TRY(cap_workspace(h, CAP_WS_STAGE0, (size_t)std::min<int64_t>(rows_per, std::max<int64_t>(m_loc, 1)) * n * sizeof(double), &T));
  CAP_TRY(cap_residual_begin(h));
  for (int64_t r0 = 0; r0 < m_loc; r0 += rows_per) {
    const int64_t rows = std::min<int64_t>(rows_per, m_loc - r0);
    CAP_TRY(trmm_impl(h, 'R', 'U', 'N', 'N', rows, n, 1.0, R, ldr, Q + r0, ldq, T, rows, true, false));
    CAP_TRY(cap_residual_accumulate(h, rows, n, T, rows, A + r0, lda, 1, 0, 0, 0, 1.0));
  }
  return cap_residual_finish(h, comm, resid);
}

struct StreamScope {   // the helpers launch on h->stream: switch it (and the split-K workspace bank) for a scope
  cap_handle_t h; cudaStream_t saved; int saved_bank;
  StreamScope(cap_handle_t h_, cudaStream_t s, int bank) : h(h_), saved(h_->stream), saved_bank(h_->ws_bank) { h->stream = s; h->ws_bank = bank; }
  ~StreamScope() { h->stream = saved; h->ws_bank = saved_bank; }
};

struct SmCapScope {
  cap_handle_t h; int saved;
  SmCapScope(cap_handle_t h_, int cap) : h(h_), saved(h_->num_sms) { if (cap > 0 && cap < saved) h->num_sms = cap; }
  ~SmCapScope() { h->num_sms = saved; }
};

#include "grid_algs.cuh"
#include "potrf_1d.cuh"

int cap_axpby2d(cap_handle_t h, double* y, int64_t ldy, const double* x, int64_t ldx, int64_t m, int64_t n, double alpha, double beta) {
  return axpby2d(h, y, ldy, x, ldx, m, n, alpha, beta, CTRI_FULL);
}

// ------------------------------------------------------------------------------------------ allocation-free entry point
// cap_cqr_gram_noalloc: G(upper) = Q^T Q like cap_cqr_gram, for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// The split-K partials that cap_cqr_gram keeps in the handle's grown-on-demand workspace live here in a
// file-scope device array of NOALLOC_SCRATCH_BYTES (32 MiB; every GPU that loads the library holds it until exit), so the
// call makes no cudaMalloc / cudaFree and calls no allocating routine.  Limits: n <= CAP_GRAM_NOALLOC_MAX_N (512).
// The array is shared by every call on the device and keeps the last call's contents: nothing in it has to start at zero
// (every partial element the reduction reads is written by the product launch of the same call), but two calls that
// overlap on the GPU (different streams or handles of one device) race on it.
constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)32 << 20;
constexpr size_t NOALLOC_SCRATCH_DOUBLES = NOALLOC_SCRATCH_BYTES / sizeof(double);
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol: its DEVICE address comes from
// cudaGetSymbolAddress, once per device, with that device current (cap_create has just made it current).
static double* g_noalloc_dev_ptr[64] = {};
static int noalloc_setup(cap_handle_s* h) {
  if (h->device < 0 || h->device >= 64) return cap_fail(CAP_ERR_ARG, "cap_create", "device index beyond 63");
  if (!g_noalloc_dev_ptr[h->device]) {
    double* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) { cudaGetLastError(); return cap_fail(CAP_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", cudaGetErrorString(e)); }
    g_noalloc_dev_ptr[h->device] = ptr;
  }
  h->noalloc_scratch = g_noalloc_dev_ptr[h->device];
  return CAP_OK;
}

int cap_cqr_gram_noalloc(cap_handle_t h, int64_t m, int64_t n, const double* Q, int64_t ldq, double* G, int64_t ldg) {
  if (m < 0 || n < 0 || m > INT32_MAX) return cap_fail(CAP_ERR_ARG, "cap_cqr_gram_noalloc", "bad dimension");
  if (n > CAP_GRAM_NOALLOC_MAX_N)
    return cap_fail(CAP_ERR_ARG, "cap_cqr_gram_noalloc", "n exceeds CAP_GRAM_NOALLOC_MAX_N (512): the static scratch holds no more; use cap_cqr_gram");
  if (ldq < std::max<int64_t>(1, m) || ldg < std::max<int64_t>(1, n)) return cap_fail(CAP_ERR_ARG, "cap_cqr_gram_noalloc", "leading dimension too small");
  if (!h->noalloc_scratch) return cap_fail(CAP_ERR_CUDA, "cap_cqr_gram_noalloc", "static scratch address was not resolved at cap_create");
  if (n == 0) return CAP_OK;
  GemmParams p{};
  p.A = Q; p.B = Q; p.lda = ldq; p.ldb = ldq; p.C = G; p.Cin = G; p.ldc = ldg; p.ldcin = ldg;
  p.M = (int)n; p.N = (int)n; p.K = (int)m; p.alpha = 1.0; p.beta = 0.0; p.c_tri = CTRI_UPPER;
  p.tiles_m = p.tiles_n = (int)((n + BM - 1) / BM);
  const int ntiles = p.tiles_m * (p.tiles_m + 1) / 2;
  // k-slabs: about two work units per SM, slabs of at least 256 rows, and no more partials than the scratch holds
  int64_t ksplit = std::min<int64_t>({(int64_t)(2 * h->num_sms) / ntiles, m / 256, (int64_t)(NOALLOC_SCRATCH_DOUBLES / (size_t)(n * n))});
  ksplit = std::max<int64_t>(1, ksplit);
  if (ksplit > 1) {
    int kper = (int)((m + ksplit - 1) / ksplit);
    kper = ((kper + BK - 1) / BK) * BK;
    ksplit = (m + kper - 1) / kper;
    if (ksplit > 1) { p.k_per_split = kper; p.c_split_stride = n * n; p.C = h->noalloc_scratch; p.ldc = n; }
  }
  p.ksplit = (int)ksplit;
  if (tma_eligible(h, p)) {
    CUtensorMap ma, mb;
    CAP_TRY(make_map(h, &ma, p.A, p.K, p.M, p.lda, true));
    CAP_TRY(make_map(h, &mb, p.B, p.K, p.N, p.ldb, true));
    TmaGemmParams tp{};
    tp.C = p.C; tp.Cin = p.Cin; tp.ldc = p.ldc; tp.ldcin = p.ldcin; tp.M = p.M; tp.N = p.N; tp.K = p.K;
    tp.alpha = 1.0; tp.beta = 0.0; tp.c_tri = CTRI_UPPER; tp.diag = DIAG_NONE;
    tp.ksplit = p.ksplit; tp.k_per_split = p.k_per_split; tp.c_split_stride = p.c_split_stride;
    tp.tiles_m = p.tiles_m; tp.tiles_n = p.tiles_n; tp.sched = SCHED_TRI; tp.subs = 1; tp.units = ntiles * p.ksplit;
    tp.c_vec2 = aligned16(tp.C, tp.ldc) ? 1 : 0;
    launch_tma_t<true, true>(h, ma, mb, tp, std::min(tp.units, h->num_sms));
  } else {
    launch_gemm_t<true, false>(h, p, 1, dim3(ntiles, p.ksplit));
  }
  LAUNCH_CHECK(h);
  if (p.ksplit > 1) {
    splitk_reduce_kernel<<<dim3((unsigned)((n + 127) / 128), (unsigned)n), 128, 0, h->stream>>>(h->noalloc_scratch, p.c_split_stride, (int)n, p.ksplit, G, ldg,
                                                                                            (int)n, (int)n, 1.0, 0.0, CTRI_UPPER);
    LAUNCH_CHECK(h);
  }
  return CAP_OK;
}
