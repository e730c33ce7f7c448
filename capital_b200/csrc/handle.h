// Internal: the opaque handle behind cap_handle_t and helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

enum { CAP_WS_GEMM_SPLITK = 0, CAP_WS_TRI = 1, CAP_WS_PANEL = 2, CAP_WS_FACTOR = 3, CAP_WS_STAGE0 = 4, CAP_WS_STAGE1 = 5, CAP_WS_GRAM = 6,
       CAP_WS_SUMMA_A = 7, CAP_WS_SUMMA_B = 8, CAP_WS_SUMMA_C = 9, CAP_WS_SUMMA_T = 10,      // broadcast landing buffers / partial C / grid-transposed copy
       CAP_WS_CI_T1 = 11, CAP_WS_CI_W = 12, CAP_WS_CI_T2 = 13, CAP_WS_CI_BC = 14, CAP_WS_CI_BC2 = 15,  // distributed cholinv intermediates
       CAP_WS_CQR3 = 16,                                                                       // 3D CQR: second Gram / R2 and the inverse factor
       CAP_WS_VALID = 17,                                                                      // validators: accumulators, block partials, product chunk
       CAP_WS_GEMM_SPLITK2 = 18,                                                               // split-K partials of launches on the panel stream (bank 1)
       CAP_WS_POTRF1D = 19,                                                                    // look-ahead Cholesky: inverse block, row panels (double-buffered)
       CAP_WS_GEMM_SPLITK3 = 20,                                                               // split-K partials of launches on the row stream (bank 2)
       CAP_WS_SLOTS = 21 };   // (CAP_WS_SUMMA_A: chunked left operand; CAP_WS_SUMMA_B: two right-operand chunk buffers)

struct cap_ws_slot { void* ptr = nullptr; size_t bytes = 0; };

// Symmetric peer-mapped buffer (comm.cu): every rank of a communicator owns `bytes` of device memory and has every other
// rank's buffer mapped through CUDA IPC, so copy engines (cudaMemcpyAsync) and kernels can write into a peer's HBM over
// NVLink / NVSwitch without any NCCL kernel resident on an SM.
enum { CAP_PEER_MAX = 16 };
struct cap_peer_s {
  int rank = 0, size = 1;
  size_t bytes = 0;
  char* ptr[CAP_PEER_MAX] = {};        // ptr[rank] = the local allocation
  struct cap_comm_s* comm = nullptr;   // the communicator it was made for (not owned)
  uint32_t seq = 0;                    // monotonic signal counter (same value on every rank: calls are collective)
};
int cap_peer_alloc(struct cap_handle_s* h, struct cap_comm_s* comm, size_t bytes, cap_peer_s** out);   // collective
int cap_peer_free(cap_peer_s* p);
int cap_peer_free_collective(struct cap_handle_s* h, cap_peer_s* p);

struct cap_handle_s {
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;      // current stream (caller's or own)
  cudaStream_t own_stream = nullptr;
  cudaStream_t copy_stream = nullptr; // host<->device staging for the e2e entry points
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;
  cudaStream_t panel_stream = nullptr; // look-ahead Cholesky: high-priority stream of the panel operations
  cudaEvent_t ev_panel[2] = {nullptr, nullptr}, ev_u1 = nullptr;
  cudaStream_t gather_stream = nullptr;  // pipelined look-ahead: stream of the row-panel all-gathers
  cudaEvent_t ev_trmm = nullptr, ev_gath[2] = {nullptr, nullptr}, ev_u2 = nullptr;
  struct cap_comm_s* dup_comm = nullptr;     // duplicate of dup_parent for those all-gathers (owned by the handle)
  struct cap_comm_s* dup_parent = nullptr;
  int lookahead_pipeline = -1;        // -1 = the asynchronous schedule (potrf_1d.cuh), 0 / 1 = round-1 two-stream / pipelined NCCL orders
  // asynchronous look-ahead Cholesky (potrf_1d.cuh): ring of events, exchange / signal streams, peer-mapped landing buffers
  enum { RING = 4 };
  cudaStream_t xchg_stream = nullptr, sig_stream = nullptr, row_stream = nullptr;
  cudaEvent_t ev_leaf[RING] = {}, ev_csolve[RING] = {}, ev_solve[RING] = {}, ev_xchg[RING] = {}, ev_xpush[RING] = {}, ev_rowready[RING] = {}, ev_r1[RING] = {}, ev_r2a[RING] = {}, ev_r2[RING] = {};
  cap_peer_s* peer1d = nullptr;       // landing buffers of the panel exchange (owned)
  int peer_exchange = -1;             // -1 = peer copy engines when CUDA IPC works, else NCCL; 0 = NCCL orders of round 1; 1 = required
  void* fn_wait32 = nullptr;          // cuStreamWaitValue32 / cuStreamWriteValue32 (driver entry points)
  void* fn_write32 = nullptr;
  uint32_t* sig_scratch = nullptr;    // one word per signalling stream: the value a 4-byte peer copy carries
  struct cap_comm_s* peer_failed_comm = nullptr;   // CUDA IPC mapping failed on this communicator: do not retry every call
  int lookahead_grid_mult = 64;       // tile-engine grid multiplier of the trailing update beside a chain (0 = keep the persistent grid)
  // chunked SUMMA (grid_algs.cuh): broadcasts / depth all-reduce on their own stream beside the local products
  // pipelined panel loop: -1 = as the reference (only when the grid's num_chunks > 0, summa.hpp:177), 0 = never, 1 = always
  int summa_pipeline = -1, summa_reserve_sms = 0, summa_halves = 1;
  cudaStream_t summa_stream = nullptr;
  cudaEvent_t ev_sready[2] = {}, ev_sfree[2] = {}, ev_shalf = nullptr, ev_sred[2] = {}, ev_sstart = nullptr;
  int cacqr_complete_inv = 1;         // 3D CholeskyQR: 1 = Q <- Q Rinv with the completed inverse, 0 = cacqr::solve (block back-substitution)
  double* noalloc_scratch = nullptr;  // device address of the static scratch of cap_cqr_gram_noalloc (cudaGetSymbolAddress at cap_create)
  int lookahead_stair = 1;            // trailing update of the asynchronous schedule as ONE staircase launch (0 = a gemm + a syrk per run of blocks)
  int grid_mult = 1;                  // tile engine: CTAs per SM worth of grid (1 = persistent, one CTA per SM)
  int ws_bank = 0;                    // which split-K workspace slot gemm launches use (1 while launching on the panel stream)
  int lookahead_reserve_sms = 16;     // SMs the trailing update leaves free while a panel runs beside it
  int potrf_schedule = 0;             // cap_dpotrf on more than one block: 0 = right-looking with look-ahead, 1 = recursive
  int* info_dev = nullptr;
  long long launches = 0;
  void* encode_tiled = nullptr;       // cuTensorMapEncodeTiled
  size_t host_chunk_bytes = (size_t)1 << 30;   // row-chunk size of the pipelined host entry points
  int engine = 0;                     // 0 = auto (TMA engine when alignment allows), 1 = cp.async engine only
  int potrf_nb = 0;                   // cap_dpotrf: largest diagonal block factored + inverted by the cholinv recursion (0 = by size)
  int trsm_nb = 512;                  // cap_dtrsm: largest diagonal block applied through its explicit inverse
  cap_ws_slot ws[CAP_WS_SLOTS];
};

int cap_fail(int code, const char* what, const char* detail);
int cap_workspace(cap_handle_s* h, int slot, size_t bytes, double** out);
