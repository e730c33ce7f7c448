// Collectives of the path over NCCL (NVLink 5 / NVSwitch), replacing the reference's MPI calls
// (Allreduce policy.h:22,82; Bcast summa.hpp:185,193; depth Allreduce summa.hpp:236;
// Sendrecv_replace util.hpp:240; Allgather cholinv/policy.h:176,240; Reduce cacqr.hpp:98).
#include "../../include/capital_b200.h"
#include "handle.h"

#include <nccl.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct cap_comm_s {
  ncclComm_t comm;
  int rank, size;
};

#define NCCL_TRY(expr)                                                                     \
  do {                                                                                     \
    ncclResult_t r_ = (expr);                                                              \
    if (r_ != ncclSuccess) return cap_fail(CAP_ERR_NCCL, #expr, ncclGetErrorString(r_));    \
  } while (0)

int cap_comm_unique_id(char* id128) {
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  NCCL_TRY(ncclGetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return CAP_OK;
}
int cap_comm_init_rank(cap_comm_t* out, const char* id128, int nranks, int rank) {
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  cap_comm_s* c = new cap_comm_s();
  c->rank = rank; c->size = nranks;
  NCCL_TRY(ncclCommInitRank(&c->comm, nranks, id, rank));
  *out = c;
  return CAP_OK;
}
int cap_comm_split(cap_comm_t parent, int color, int key, cap_comm_t* child) {
  cap_comm_s* c = new cap_comm_s();
  NCCL_TRY(ncclCommSplit(parent->comm, color, key, &c->comm, nullptr));
  NCCL_TRY(ncclCommUserRank(c->comm, &c->rank));
  NCCL_TRY(ncclCommCount(c->comm, &c->size));
  *child = c;
  return CAP_OK;
}
int cap_comm_destroy(cap_comm_t c) {
  if (!c) return CAP_OK;
  ncclCommDestroy(c->comm);
  delete c;
  return CAP_OK;
}
int cap_comm_rank(cap_comm_t c, int* r) { *r = c->rank; return CAP_OK; }
int cap_comm_size(cap_comm_t c, int* s) { *s = c->size; return CAP_OK; }

int cap_allreduce_sum(cap_handle_t h, cap_comm_t c, double* buf, int64_t count) {
  if (!c || c->size == 1 || count == 0) return CAP_OK;
  NCCL_TRY(ncclAllReduce(buf, buf, (size_t)count, ncclDouble, ncclSum, c->comm, h->stream));
  return CAP_OK;
}
int cap_reduce_sum(cap_handle_t h, cap_comm_t c, double* buf, int64_t count, int root) {
  if (!c || c->size == 1 || count == 0) return CAP_OK;
  NCCL_TRY(ncclReduce(buf, buf, (size_t)count, ncclDouble, ncclSum, root, c->comm, h->stream));
  return CAP_OK;
}
int cap_bcast(cap_handle_t h, cap_comm_t c, double* buf, int64_t count, int root) {
  if (!c || c->size == 1 || count == 0) return CAP_OK;
  NCCL_TRY(ncclBroadcast(buf, buf, (size_t)count, ncclDouble, root, c->comm, h->stream));
  return CAP_OK;
}
int cap_allgather(cap_handle_t h, cap_comm_t c, const double* send, double* recv, int64_t count) {
  if (!c || c->size == 1) {
    if (send != recv && count) { if (cudaMemcpyAsync(recv, send, (size_t)count * 8, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess) return cap_fail(CAP_ERR_CUDA, "cap_allgather", "copy failed"); }
    return CAP_OK;
  }
  NCCL_TRY(ncclAllGather(send, recv, (size_t)count, ncclDouble, c->comm, h->stream));
  return CAP_OK;
}
// pairwise exchange (the grid transpose): NCCL has no in-place Sendrecv_replace, so send and recv differ
int cap_sendrecv(cap_handle_t h, cap_comm_t c, const double* send, double* recv, int64_t count, int peer) {
  if (!c || peer == c->rank) {
    if (send != recv && count) { if (cudaMemcpyAsync(recv, send, (size_t)count * 8, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess) return cap_fail(CAP_ERR_CUDA, "cap_sendrecv", "copy failed"); }
    return CAP_OK;
  }
  NCCL_TRY(ncclGroupStart());
  NCCL_TRY(ncclSend(send, (size_t)count, ncclDouble, peer, c->comm, h->stream));
  NCCL_TRY(ncclRecv(recv, (size_t)count, ncclDouble, peer, c->comm, h->stream));
  NCCL_TRY(ncclGroupEnd());
  return CAP_OK;
}

// ------------------------------------------------------------------------------------------ peer-mapped buffers
// cap_peer_alloc: collective over `comm`.  Every rank cudaMallocs `bytes` (zeroed), the 64-byte CUDA IPC handles are
// all-gathered through NCCL (the only NCCL call of this path), and every other rank's buffer is opened with
// cudaIpcOpenMemHandle, which also enables peer access over NVLink.  If ANY rank cannot map a peer, every rank gets
// CAP_ERR_UNSUPPORTED (agreed by a second tiny collective) and the caller keeps its NCCL path.
int cap_peer_alloc(cap_handle_s* h, cap_comm_s* comm, size_t bytes, cap_peer_s** out) {
  *out = nullptr;
  if (!comm || comm->size < 2 || comm->size > CAP_PEER_MAX) return cap_fail(CAP_ERR_UNSUPPORTED, "cap_peer_alloc", "needs 2..16 ranks");
  const int P = comm->size, me = comm->rank;
  cap_peer_s* p = new cap_peer_s();
  p->rank = me; p->size = P; p->bytes = bytes; p->comm = comm;
  int bad = 0;
  void* local = nullptr;
  std::string why;
  auto note = [&](const char* step, cudaError_t e) { why = std::string(step) + ": " + cudaGetErrorString(e); cudaGetLastError(); bad = 1; };
  cudaError_t ce;
  if ((ce = cudaMalloc(&local, bytes)) != cudaSuccess) note("cudaMalloc", ce);
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (!bad && (ce = cudaMemset(local, 0, bytes)) != cudaSuccess) note("cudaMemset", ce);
  if (!bad && (ce = cudaIpcGetMemHandle(&mine, local)) != cudaSuccess) note("cudaIpcGetMemHandle", ce);
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  char* stage = nullptr;                                   // [P handles][P status words]
  const size_t hb = sizeof(cudaIpcMemHandle_t);
  if (cudaMalloc((void**)&stage, (size_t)P * hb + (size_t)P * hb) != cudaSuccess) { cudaGetLastError(); delete p; if (local) cudaFree(local); return cap_fail(CAP_ERR_ALLOC, "cap_peer_alloc", "staging"); }
  std::vector<cudaIpcMemHandle_t> all(P);
  cudaMemcpy(stage + (size_t)me * hb, &mine, hb, cudaMemcpyHostToDevice);
  cudaStreamSynchronize(h->stream);
  ncclResult_t nr = ncclAllGather(stage + (size_t)me * hb, stage, hb, ncclChar, comm->comm, h->stream);
  if (nr != ncclSuccess || cudaStreamSynchronize(h->stream) != cudaSuccess) { cudaFree(stage); if (local) cudaFree(local); delete p; return cap_fail(CAP_ERR_NCCL, "cap_peer_alloc", "handle exchange failed"); }
  cudaMemcpy(all.data(), stage, (size_t)P * hb, cudaMemcpyDeviceToHost);
  p->ptr[me] = (char*)local;
  for (int r = 0; r < P && !bad; ++r) {
    if (r == me) continue;
    void* q = nullptr;
    if ((ce = cudaIpcOpenMemHandle(&q, all[r], cudaIpcMemLazyEnablePeerAccess)) != cudaSuccess) { note("cudaIpcOpenMemHandle", ce); break; }
    p->ptr[r] = (char*)q;
  }
  // agree on the outcome: sum of the per-rank failure flags
  float flag = (float)bad, total = 0.f;
  float* dflag = (float*)(stage + (size_t)P * hb);
  cudaMemcpy(dflag, &flag, sizeof(float), cudaMemcpyHostToDevice);
  nr = ncclAllReduce(dflag, dflag, 1, ncclFloat, ncclSum, comm->comm, h->stream);
  cudaStreamSynchronize(h->stream);
  cudaMemcpy(&total, dflag, sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(stage);
  if (nr != ncclSuccess || total != 0.f) {
    cap_peer_free(p);
    if (getenv("CAPITAL_B200_VERBOSE") || bad) fprintf(stderr, "[capital_b200] rank %d: peer mapping of %zu bytes unavailable (%s); NCCL exchange is used instead\n", me, bytes, bad ? why.c_str() : "another rank failed");
    return cap_fail(CAP_ERR_UNSUPPORTED, "cap_peer_alloc", bad ? why.c_str() : "CUDA IPC peer mapping failed on another rank");
  }
  *out = p;
  return CAP_OK;
}

// Collective variant (re-allocation in mid-job): every rank closes its mappings, all meet, then the exporters free.
int cap_peer_free_collective(cap_handle_s* h, cap_peer_s* p) {
  if (!p) return CAP_OK;
  cudaDeviceSynchronize();
  for (int r = 0; r < p->size; ++r)
    if (r != p->rank && p->ptr[r]) { cudaIpcCloseMemHandle(p->ptr[r]); p->ptr[r] = nullptr; }
  float* d = nullptr;
  if (p->comm && cudaMalloc((void**)&d, sizeof(float)) == cudaSuccess) {
    cudaMemset(d, 0, sizeof(float));
    ncclAllReduce(d, d, 1, ncclFloat, ncclSum, p->comm->comm, h->stream);
    cudaStreamSynchronize(h->stream);
    cudaFree(d);
  }
  return cap_peer_free(p);
}

// Not collective: the caller makes sure (device synchronised, a barrier or the end of the job) that no peer still writes here.
int cap_peer_free(cap_peer_s* p) {
  if (!p) return CAP_OK;
  for (int r = 0; r < p->size; ++r) {
    if (!p->ptr[r]) continue;
    if (r == p->rank) cudaFree(p->ptr[r]); else cudaIpcCloseMemHandle(p->ptr[r]);
  }
  cudaGetLastError();
  delete p;
  return CAP_OK;
}
