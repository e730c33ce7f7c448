/* capital_b200 -- C ABI of the B200-native factorization hot path.
 *
 * Drop-in boundary for huttered40/capital's engine layer (reference paths relative to the repo root):
 *   blas::engine::_gemm/_trmm/_syrk        src/blas/interface.h:57-66   (double specialisations interface.hpp:43-97)
 *   lapack::engine::_potrf/_trtri          src/lapack/interface.h:49-53 (interface.hpp:31-58)
 * plus the memory/serialize, generator and collective calls the algorithms make around them, so
 * that the schedules of src/alg/{qr/cacqr,cholesky/cholinv,matmult/summa} run on device buffers.
 *
 * Conventions: every matrix pointer is a DEVICE pointer on the handle's device, column-major, FP64;
 * calls are stream-ordered on the handle's stream and do not synchronise unless stated; flags are
 * the BLAS characters ('N','T' / 'U','L' / 'L','R' / 'N','U'); return value 0 = success, otherwise a
 * CAP_* status (never an exception).  potrf/trtri report LAPACK-style `info` through a device int
 * owned by the handle (cap_get_info synchronises and returns it).  There is no CPU fallback: if no
 * CUDA device is usable, cap_create fails.
 */
#ifndef CAPITAL_B200_H_
#define CAPITAL_B200_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct cap_handle_s* cap_handle_t;
typedef struct cap_comm_s* cap_comm_t;

enum { CAP_OK = 0, CAP_ERR_ARG = 1, CAP_ERR_CUDA = 2, CAP_ERR_ALLOC = 3, CAP_ERR_NCCL = 4, CAP_ERR_UNSUPPORTED = 5 };

/* ---- context ------------------------------------------------------------------------------- */
int cap_create(cap_handle_t* h, int device);
int cap_destroy(cap_handle_t h);
int cap_set_stream(cap_handle_t h, void* cuda_stream);      /* cudaStream_t; NULL = the legacy default stream */
int cap_use_own_stream(cap_handle_t h);                     /* back to the handle's private non-blocking stream (the default) */
void* cap_get_stream(cap_handle_t h);
int cap_synchronize(cap_handle_t h);
const char* cap_last_error(void);
int cap_get_info(cap_handle_t h, int* info);                 /* syncs; LAPACK info of potrf/trtri since last reset */
int cap_reset_info(cap_handle_t h);
long long cap_kernel_launches(cap_handle_t h);               /* number of kernels this handle has launched */
/* tile engine selection: 0 = auto (TMA/mbarrier warp-specialised engine whenever operands are 16-byte aligned
 * with even leading dimensions, cp.async engine otherwise), 1 = cp.async engine only (A/B testing) */
int cap_set_engine(cap_handle_t h, int engine);
/* row-chunk size (bytes, default 1 GiB) of the pipelined host entry point cap_cacqr_factor_1d_host */
int cap_set_host_chunk_bytes(cap_handle_t h, size_t bytes);
/* synchronise the device and free every library-owned workspace slot (they are grown on demand and otherwise kept
 * until cap_destroy); a caller that moves from one large problem to another uses this to return the HBM */
int cap_release_workspace(cap_handle_t h);

/* ---- memory (matrix<>::_assemble / _destroy_, src/matrix/structure.hpp:4-24, matrix.hpp:158) -- */
int cap_malloc(void** dptr, size_t bytes);
int cap_free(void* dptr);
int cap_memset_async(cap_handle_t h, void* dptr, int value, size_t bytes);
int cap_h2d_async(cap_handle_t h, void* dst, const void* src, size_t bytes);
int cap_d2h_async(cap_handle_t h, void* dst, const void* src, size_t bytes);
int cap_host_alloc(void** hptr, size_t bytes);               /* pinned */
int cap_host_free(void* hptr);

/* ---- BLAS-3 engine (src/blas/interface.hpp:43-97) -------------------------------------------- */
int cap_dgemm(cap_handle_t h, char transa, char transb, int64_t m, int64_t n, int64_t k, double alpha,
              const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc);
int cap_dsyrk(cap_handle_t h, char uplo, char trans, int64_t n, int64_t k, double alpha,
              const double* A, int64_t lda, double beta, double* C, int64_t ldc);
/* in place, B <- alpha*op(T)*B (side 'L') or alpha*B*op(T) (side 'R'); only `uplo` triangle of T is read */
int cap_dtrmm(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
              const double* T, int64_t ldt, double* B, int64_t ldb);
/* out of place, C <- alpha*op(T)*B or alpha*B*op(T); B and C must not overlap */
int cap_dtrmm_oop(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
                  const double* T, int64_t ldt, const double* B, int64_t ldb, double* C, int64_t ldc);

/* triangular solve, in place: op(T) X = alpha*B (side 'L') or X op(T) = alpha*B (side 'R'), non-unit diagonal.  The
 * reference's engine has no _trsm (src/blas/interface.h:57-66) -- its schedules multiply by explicit inverses
 * (cholinv.hpp:113-124, the unfinished src/alg/trsm/diaginvert); this entry point is the solve those stand for, and the
 * building block that makes cap_dpotrf an N^3/3 schedule: recursion on T, off-diagonal updates as plain gemms on the
 * tile engine, diagonal blocks of at most trsm_nb applied through their inverses. */
int cap_dtrsm(cap_handle_t h, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
              const double* T, int64_t ldt, double* B, int64_t ldb);
/* blocking of cap_dpotrf (largest diagonal block factored + inverted by the cholinv recursion; default 0 = by size:
 * n/8 clamped to [512, 2048]) and cap_dtrsm (largest diagonal block applied through its inverse, default 512);
 * multiples of 128 */
int cap_set_block_sizes(cap_handle_t h, int potrf_nb, int trsm_nb);

/* ---- LAPACK engine (src/lapack/interface.hpp:31-58) ------------------------------------------ */
/* 'U' only (cholinv.hpp:9).  N^3/3 flops: recursive potrf / trsm / triangular syrk; only diagonal blocks of at most
 * potrf_nb are inverted (cap_dcholinv below is the reference's schedule, which inverts at every level). */
int cap_dpotrf(cap_handle_t h, char uplo, int n, double* A, int lda);
int cap_dtrtri(cap_handle_t h, char uplo, char diag, int n, double* A, int lda);  /* 'U' or 'L', non-unit */
/* fused potrf + trtri as the reference always uses them (cacqr.hpp:20-22, cholinv/policy.h:199-201):
 * A(upper) <- R, Rinv(upper) <- R^-1, strict lower of Rinv zeroed.  complete_inv=0 leaves the top-level
 * off-diagonal block of Rinv unset, exactly like cholinv (cholinv.hpp:147). */
int cap_dcholinv(cap_handle_t h, int n, double* A, int64_t lda, double* Rinv, int64_t ldri, int complete_inv);

/* ---- CholeskyQR fast paths (cacqr.hpp:7-29) -------------------------------------------------- */
/* G(upper, n x n, ldg) <- Q^T Q for the local m x n row block; deterministic split-K. */
int cap_cqr_gram(cap_handle_t h, int64_t m, int64_t n, const double* Q, int64_t ldq, double* G, int64_t ldg);
/* Same result as cap_cqr_gram (upper triangle of G <- Q^T Q; summation order of the k-slabs may differ) WITHOUT any
 * allocation: the split-K partials live in a 32 MiB file-scope device array instead of the handle's workspace, so the
 * call is usable under CUDA graph capture or wherever cudaMalloc is forbidden.  No counterpart in the reference.
 * Limit: n <= CAP_GRAM_NOALLOC_MAX_N, larger n returns CAP_ERR_ARG (any m).
 * Every call on a device shares that array: calls that overlap on the GPU (two streams or handles of one device) race. */
#define CAP_GRAM_NOALLOC_MAX_N 512
int cap_cqr_gram_noalloc(cap_handle_t h, int64_t m, int64_t n, const double* Q, int64_t ldq, double* G, int64_t ldg);
/* Qout <- Qin * triu(Rinv); Qout may equal Qin (in place) or be a different buffer. */
int cap_cqr_apply(cap_handle_t h, int64_t m, int64_t n, const double* Rinv, int64_t ldr,
                  const double* Qin, int64_t ldqin, double* Qout, int64_t ldqout);

/* ---- serialize / util (src/matrix/serialize.hpp:12-150, src/util/util.hpp:267-318, summa.hpp:33) */
/* mode: 0 all, 1 upper only, 2 lower only, 3 upper + zero strict lower, 4 lower + zero strict upper */
int cap_copy2d(cap_handle_t h, double* dst, int64_t ldd, const double* src, int64_t lds, int64_t m, int64_t n, int mode);
int cap_pack_upper(cap_handle_t h, double* packed, const double* full, int64_t ld, int n);
int cap_unpack_upper(cap_handle_t h, double* full, int64_t ld, const double* packed, int n, int zero_lower);
int cap_axpby(cap_handle_t h, double* y, const double* x, int64_t count, double alpha, double beta);
/* y(m x n, ldy) = alpha * x(m x n, ldx) + beta * y: the strided form (views of a larger local block) */
int cap_axpby2d(cap_handle_t h, double* y, int64_t ldy, const double* x, int64_t ldx, int64_t m, int64_t n, double alpha, double beta);
/* dst[i + i*ld] = value for i < n (rect::_distribute_identity, src/matrix/structure.hpp:37-66, on a zeroed block) */
int cap_fill_diagonal(cap_handle_t h, double* dst, int64_t ld, int64_t n, double value);

/* ---- generators (src/matrix/structure.hpp:69-129), bit-exact with glibc drand48 -------------- */
int cap_fill_random(cap_handle_t h, double* dst, int64_t ld, int64_t rows, int64_t cols, long long key);
int cap_fill_symmetric(cap_handle_t h, double* dst, int64_t ld, int64_t rows, int64_t cols, int64_t px, int64_t py,
                       int64_t grid_x, int64_t grid_y, int64_t global_dim, int diagonally_dominant);

/* ---- collectives: NCCL in place of the reference's MPI ------------------------ */
int cap_comm_unique_id(char* id128);                          /* 128-byte ncclUniqueId */
int cap_comm_init_rank(cap_comm_t* comm, const char* id128, int nranks, int rank);
int cap_comm_split(cap_comm_t parent, int color, int key, cap_comm_t* child);
int cap_comm_destroy(cap_comm_t comm);
int cap_comm_rank(cap_comm_t comm, int* rank);
int cap_comm_size(cap_comm_t comm, int* size);
int cap_allreduce_sum(cap_handle_t h, cap_comm_t comm, double* buf, int64_t count);              /* in place */
int cap_reduce_sum(cap_handle_t h, cap_comm_t comm, double* buf, int64_t count, int root);        /* in place at root */
int cap_bcast(cap_handle_t h, cap_comm_t comm, double* buf, int64_t count, int root);
int cap_allgather(cap_handle_t h, cap_comm_t comm, const double* send, double* recv, int64_t count_per_rank);
int cap_sendrecv(cap_handle_t h, cap_comm_t comm, const double* send, double* recv, int64_t count, int peer);

/* ---- process grid: topo::square (src/util/topology.h:67-143) and the algorithms that run on it ---------------
 * c x d x d grid over `world` (size == c*d*d, d % c == 0 or d == 1), rank -> (x, y, z) as the reference's layout 0;
 * row / column / depth / slice communicators are NCCL splits.  A global matrix is dealt element-cyclically over
 * the d x d face (process (x,y) holds A[y + d*i, x + d*j]) and replicated over depth, as in src/matrix/structure.hpp:74-98.
 * All dimensions below are LOCAL block dimensions; all calls are collective over the grid and stream-ordered. */
typedef struct cap_grid_s* cap_grid_t;
int cap_grid_create(cap_comm_t world, int c, cap_grid_t* grid);      /* world may be NULL for a single rank */
/* topo::square(comm, c, layout, num_chunks) (src/util/topology.h:69): layout 0 or 1 = the two rank -> (x, y, z) orders of the
 * reference; num_chunks = pieces per panel of the pipelined SUMMA (0 = chosen from the contraction length) */
int cap_grid_create2(cap_comm_t world, int c, int layout, int num_chunks, cap_grid_t* grid);
int cap_grid_set_chunks(cap_grid_t grid, int num_chunks);
int cap_grid_destroy(cap_grid_t grid);
int cap_grid_info(cap_grid_t grid, int* c, int* d, int* x, int* y, int* z);
/* util::transpose (src/util/util.hpp:233-247): dst <- the block of the process at (y, x) in the same layer */
int cap_grid_transpose(cap_handle_t h, cap_grid_t grid, const double* src, double* dst, int64_t count);
/* matmult::summa::invoke, gemm overload (src/alg/matmult/summa/summa.hpp:7-44) */
int cap_summa_gemm(cap_handle_t h, cap_grid_t grid, char transa, char transb, int64_t m, int64_t n, int64_t k, double alpha,
                   const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc);
/* trmm overload (summa.hpp:47-83): B <- alpha*op(T)*B or alpha*B*op(T), in place */
int cap_summa_trmm(cap_handle_t h, cap_grid_t grid, char side, char uplo, char trans, char diag, int64_t m, int64_t n, double alpha,
                   const double* T, int64_t ldt, double* B, int64_t ldb);
/* syrk overloads (summa.hpp:86-161): C(uplo) <- alpha*A^T*A + beta*C ('T', A k x n) or alpha*A*A^T ('N', A n x k);
 * At = grid-transposed copy of A if the caller has one (the 4-argument overload), else NULL */
int cap_summa_syrk(cap_handle_t h, cap_grid_t grid, char uplo, char trans, int64_t n, int64_t k, double alpha,
                   const double* A, int64_t lda, const double* At, int64_t ldat, double beta, double* C, int64_t ldc);
/* cholesky::cholinv::factor (src/alg/cholesky/cholinv/cholinv.hpp:6-29): A (local n_loc x n_loc block of the SPD
 * matrix) <- local part of R (upper), Rinv <- local part of R^-1; base case = ReplicateComp (policy.h:226-305) at
 * global size <= bc_global; split as in cholinv.hpp:107 (1 = halves). */
int cap_cholinv_factor(cap_handle_t h, cap_grid_t grid, int64_t n_loc, double* A, int64_t lda, double* Rinv, int64_t ldri,
                       int complete_inv, int split, int64_t bc_global);

/* ---- Cholesky with look-ahead, one GPU or block columns dealt 1-D over the ranks of comm -----------------------------
 * A = R^T R (upper) of a global N x N SPD matrix, N^3/3 flops: what lapack::engine::_potrf computes
 * (src/lapack/interface.hpp:31-43), as a right-looking blocked schedule whose diagonal-block work (factor + inverse of
 * an nb x nb block, broadcast, row-panel solve, all-gather) runs on a second stream WHILE the previous panel's
 * trailing update occupies the tensor pipes (BASELINE.json north_star allows a 1D grid for the Cholesky target).
 * Layout: block columns of width nb (multiple of 128); block j -> round r = j / P, rank (r even ? j % P : P-1 - j % P),
 * local block r; a rank's blocks sit side by side in one column-major N x ncols_loc array (lda >= N).  Only the upper
 * triangle of the global matrix is read; on exit it holds R, everything below the diagonal is untouched.
 * comm == NULL: one rank (this is what cap_dpotrf runs when its schedule is 0). */
int cap_potrf_1d(cap_handle_t h, cap_comm_t comm, int64_t N, int nb, double* A_loc, int64_t lda);
int cap_potrf_1d_layout(int64_t N, int nb, int nranks, int rank, int64_t* ncols_loc, int* nblocks_loc);
int cap_potrf_1d_block(int64_t j, int nranks, int* owner, int* local_block);
/* test hook (host only, no device work): the tile enumeration of the one-launch trailing update of cap_potrf_1d.  Geometry:
 * T = nb / 128 tiles per block, P ranks, this rank `me`, first local block r_first, first row block row0, tiles_n column
 * tiles.  Returns the number of work units (-1: a block column left of row0); unit >= 0 also yields that unit's tile. */
int64_t cap_debug_stair_decode(int T, int P, int me, int r_first, int row0, int tiles_n, int64_t unit, int* tm, int* tn, int* tri);
/* reserve_sms: SMs the trailing update leaves to a diagonal-block chain running beside it (default 16; 0 = none).
 * pipeline: -1 (default) = the asynchronous schedule: the chain leaf -> X -> next diagonal block runs two panels ahead of
 * the trailing update on its owner, and between ranks every solved block is written into the peers' row-panel buffers by
 * the copy engines over NVSwitch (CUDA-IPC peer memory, stream memory-op flags; no NCCL kernel on any SM).  Falls back to
 * order 0 when CUDA IPC is not available between the ranks.  0 / 1 = the round-1 orders over NCCL (panel operations and
 * all-gather in one stream / the all-gather on its own stream and a duplicate communicator).
 * NOTE: the handle keeps peer mappings (and with pipeline = 1 a duplicate of comm): destroy the handle before the
 * communicator it was used with. */
int cap_set_lookahead(cap_handle_t h, int reserve_sms, int pipeline);
/* panel exchange of cap_potrf_1d: -1 (default) = peer copy engines when CUDA IPC works, else NCCL; 0 = NCCL only; 1 = required */
int cap_set_peer_exchange(cap_handle_t h, int mode);
/* tile-engine grid of the trailing update that runs beside a chain: 0 (default) = persistent, one CTA per SM, SM-capped by
 * reserve_sms; k > 0 = k x #SMs CTAs, no cap (high-priority chain kernels then get an SM at the next tile boundary) */
int cap_set_update_grid(cap_handle_t h, int ctas_per_sm);
/* cap_dpotrf on more than one potrf_nb block: 0 (default) = cap_potrf_1d on one rank, 1 = recursive potrf / trsm / syrk */
int cap_set_potrf_schedule(cap_handle_t h, int schedule);

/* ---- algorithm entry points (device-resident; src/alg/...) ----------------------------------- */
/* qr::cacqr::factor, 1D (cacqr.hpp:219-248 -> invoke_1d :174-193).  A is the local m_loc x n row block
 * (untouched), Q the output block, R the n x n upper factor (full layout, strict lower zero).
 * comm may be NULL for a single rank.  num_iter: 1 = CholeskyQR, 2 = CholeskyQR2. */
int cap_cacqr_factor_1d(cap_handle_t h, cap_comm_t comm, int64_t m_loc, int64_t n, const double* A, int64_t lda,
                        double* Q, int64_t ldq, double* R, int64_t ldr, int num_iter);
/* Same, end to end from HOST memory, pipelined over row chunks on two streams: the H2D copy of A overlaps the
 * first Gram pass, the last apply overlaps the D2H copy of Q.  Synchronises before returning.  Buffers should be
 * pinned (cap_host_alloc) for full PCIe speed. */
int cap_cacqr_factor_1d_host(cap_handle_t h, cap_comm_t comm, int64_t m_loc, int64_t n, const double* A_host, int64_t lda,
                             double* Q_host, int64_t ldq, double* R_host, int64_t ldr, int num_iter);
/* qr::cacqr::factor on a process grid: the 3D / tunable CA-CQR2 (cacqr.hpp:219-248 -> invoke_3d :197-216 ->
 * sweep_3d :77-122; grid of topo::square / topo::rect, topology.h:16-143).  A, Q: local m_loc x n_loc blocks,
 * element-cyclic over the d x d face (rows over y, columns over x), replicated over depth.  R: local n_loc x n_loc
 * block of the upper factor; Rinv (may be NULL): same for the inverse factor of the LAST sweep.  Gram by SUMMA-syrk,
 * R and R^-1 by cap_cholinv_factor (split, bc_global as there), Q <- Q R^-1 by SUMMA-trmm; always completes the
 * inverse (the reference's complete_inv == 0 back-substitution yields the same Q). */
int cap_cacqr_factor_3d(cap_handle_t h, cap_grid_t grid, int64_t m_loc, int64_t n_loc, const double* A, int64_t lda,
                        double* Q, int64_t ldq, double* R, int64_t ldr, double* Rinv, int64_t ldri,
                        int num_iter, int split, int64_t bc_global);
/* complete_inv of the CholeskyQR's cholinv argument pack (cacqr.hpp:104-110): 1 (default) = Q <- Q Rinv with the completed
 * inverse, 0 = cacqr::solve, block back-substitution with the inverse left without its top-level off-diagonal block */
int cap_set_cacqr_complete_inv(cap_handle_t h, int complete_inv);

/* ---- validators: the caller side of the path, on the device (test/qr/validate.hpp, test/cholesky/validate.hpp) ----
 * util::residual_local (src/util/util.hpp:27-53) as begin / accumulate / finish: accumulate adds sum w (X - Ref)^2 and
 * sum w Ref^2 over a local m x n block whose element (r, c) is global (y + d r, x + d c); Ref == NULL means the identity
 * (control 1 per element); upper_only keeps global row <= column; w = offdiag_weight off the global diagonal.
 * finish sums over comm (may be NULL), synchronises and returns sqrt(err) / sqrt(control) in *out (host). */
int cap_residual_begin(cap_handle_t h);
int cap_residual_accumulate(cap_handle_t h, int64_t m, int64_t n, const double* X, int64_t ldx, const double* Ref, int64_t ldref,
                            int d, int x, int y, int upper_only, double offdiag_weight);
int cap_residual_finish(cap_handle_t h, cap_comm_t comm, double* out);
/* qr::validate::orthogonality and ::residual (test/qr/validate.hpp:7-62): *orth = ||Q^T Q - I||_F / n,
 * *resid = ||Q R - A||_F / ||A||_F (host doubles); 1D row blocks over comm, or local blocks on a grid */
int cap_validate_cacqr_1d(cap_handle_t h, cap_comm_t comm, int64_t m_loc, int64_t n, const double* A, int64_t lda,
                          const double* Q, int64_t ldq, const double* R, int64_t ldr, double* orth, double* resid);
int cap_validate_cacqr_3d(cap_handle_t h, cap_grid_t grid, int64_t m_loc, int64_t n_loc, const double* A, int64_t lda,
                          const double* Q, int64_t ldq, const double* R, int64_t ldr, double* orth, double* resid);
/* cholesky::validate::residual (test/cholesky/validate.hpp:7-30): ||R^T R - A||_F / ||A||_F, upper triangle;
 * A = the original local block (cap_cholinv_factor overwrites its argument: keep a copy) */
int cap_validate_cholinv(cap_handle_t h, cap_grid_t grid, int64_t n_loc, const double* A, int64_t lda, const double* R, int64_t ldr, double* resid);

#ifdef __cplusplus
}
#endif
#endif /* CAPITAL_B200_H_ */
