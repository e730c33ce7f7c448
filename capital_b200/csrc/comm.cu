// Collectives of the path over NCCL (NVLink 5 / NVSwitch), replacing the reference's MPI calls
// (Allreduce policy.h:22,82; Bcast summa.hpp:185,193; depth Allreduce summa.hpp:236;
// Sendrecv_replace util.hpp:240; Allgather cholinv/policy.h:176,240; Reduce cacqr.hpp:98).
#include "../../include/capital_b200.h"
#include "handle.h"

#include <nccl.h>
#include <cstring>

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
