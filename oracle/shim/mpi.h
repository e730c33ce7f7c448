/* TEST INFRASTRUCTURE (oracle).  Not shipped, not on the product path.
 *
 * A miniature MPI (fork + shared memory, see minimpi.cpp) exposing exactly the entry
 * points the reference's live code uses (list: SURVEY.md section 8(c)), with MPICH-style
 * integer handles because the reference stores MPI_DOUBLE in a size_t
 * (src/util/shared.h:43,48).  A datatype handle is the element size in bytes.
 */
#pragma once
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_COMM_NULL (-1)
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_IN_PLACE ((void*)1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_THREAD_SINGLE 0
#define MPI_SUCCESS 0

int MPI_Init_thread(int* argc, char*** argv, int required, int* provided);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm c, int* rank);
int MPI_Comm_size(MPI_Comm c, int* size);
int MPI_Comm_split(MPI_Comm c, int color, int key, MPI_Comm* out);
int MPI_Comm_dup(MPI_Comm c, MPI_Comm* out);
int MPI_Comm_free(MPI_Comm* c);
int MPI_Barrier(MPI_Comm c);
double MPI_Wtime(void);
int MPI_Bcast(void* buf, long count, MPI_Datatype t, int root, MPI_Comm c);
int MPI_Ibcast(void* buf, long count, MPI_Datatype t, int root, MPI_Comm c, MPI_Request* r);
int MPI_Allreduce(const void* s, void* r, long count, MPI_Datatype t, MPI_Op op, MPI_Comm c);
int MPI_Iallreduce(const void* s, void* r, long count, MPI_Datatype t, MPI_Op op, MPI_Comm c, MPI_Request* q);
int MPI_Reduce(const void* s, void* r, long count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c);
int MPI_Allgather(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, MPI_Comm c);
int MPI_Gather(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, int root, MPI_Comm c);
int MPI_Scatter(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, int root, MPI_Comm c);
int MPI_Iscatter(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request* q);
int MPI_Alltoall(const void* s, long sc, MPI_Datatype st, void* r, long rc, MPI_Datatype rt, MPI_Comm c);
int MPI_Sendrecv_replace(void* buf, long count, MPI_Datatype t, int dest, int stag, int src, int rtag, MPI_Comm c, MPI_Status* st);
int MPI_Wait(MPI_Request* r, MPI_Status* s);
int PMPI_Barrier(MPI_Comm c);
int PMPI_Allreduce(const void* s, void* r, long count, MPI_Datatype t, MPI_Op op, MPI_Comm c);

/* launcher hook: minimpi.cpp owns the real main(), forks MINIMPI_NP ranks and calls this */
int minimpi_rank_main(int argc, char** argv);
#ifdef __cplusplus
}
#endif
#ifndef MINIMPI_IMPLEMENTATION
#define main minimpi_rank_main
#endif
