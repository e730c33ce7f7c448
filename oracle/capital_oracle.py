"""TEST INFRASTRUCTURE (oracle).  Not shipped, never imported by the product path.

CPU restatement (numpy) of the reference's factorization hot path, each function citing the
reference file:line it follows (paths relative to /root/reference).  Only tests/, bench.py's
cpu_baseline leg and __graft_entry__.smoke() may import this module.

Pinning: the reference has no golden vectors of its own; this restatement is
pinned against outputs of the UNMODIFIED reference compiled here (oracle/_ref/ref_driver, OpenBLAS
0.3.15 behind the mkl.h shim, fork/shm mini-MPI) -- fixtures under tests/golden/ written by
tests/golden/make_golden.py, checked by tests/test_oracle.py.

Conventions (reference src/matrix/structure.h:13): local storage is column-major with leading
dimension = local rows; "X" is a column index and "Y" a row index.  Arrays here are numpy arrays of
shape (rows, cols) in Fortran order, so arr[row, col] is the reference's data[col*rows + row].
"""
import numpy as np

# ----------------------------------------------------------------------------- drand48
_A = 0x5DEECE66D
_C = 0xB
_MASK = (1 << 48) - 1


def lcg_seed(seed):
    """srand48(seed): X0 = (low 32 bits of seed) << 16 | 0x330E (glibc; used at structure.hpp:72,80,84,108)."""
    return (((int(seed) & 0xFFFFFFFF) << 16) | 0x330E) & _MASK


def _affine_pow(t):
    """(a_t, c_t) with X_{k+t} = a_t X_k + c_t mod 2^48 (jump-ahead by repeated squaring)."""
    a_t, c_t = 1, 0
    a, c = _A, _C
    while t:
        if t & 1:
            a_t, c_t = (a * a_t) & _MASK, (a * c_t + c) & _MASK
        a, c = (a * a) & _MASK, (a * c + c) & _MASK
        t >>= 1
    return a_t, c_t


def drand48_stream(seed, count):
    """First `count` values of drand48() after srand48(seed): value_k = X_{k+1} / 2^48."""
    out = np.empty(count, dtype=np.float64)
    if count == 0:
        return out
    # vectorised in blocks: state_{k+1..k+B} from state_k through precomputed affine maps
    B = min(count, 1 << 16)
    amul = np.empty(B, dtype=object)
    cadd = np.empty(B, dtype=object)
    a_t, c_t = 1, 0
    for i in range(B):
        a_t, c_t = (_A * a_t) & _MASK, (_A * c_t + _C) & _MASK
        amul[i], cadd[i] = a_t, c_t
    # split 48-bit multiplies into 24-bit limbs so uint64 arithmetic cannot overflow
    a_lo = np.array([int(v) & 0xFFFFFF for v in amul], dtype=np.uint64)
    a_hi = np.array([int(v) >> 24 for v in amul], dtype=np.uint64)
    cc = np.array([int(v) for v in cadd], dtype=np.uint64)
    x = lcg_seed(seed)
    pos = 0
    m48 = np.uint64(_MASK)
    while pos < count:
        nb = min(B, count - pos)
        x_lo = np.uint64(x & 0xFFFFFF)
        x_hi = np.uint64(x >> 24)
        lo = a_lo[:nb] * x_lo
        mid = (a_lo[:nb] * x_hi + a_hi[:nb] * x_lo) & np.uint64(0xFFFFFF)
        st = (lo + (mid << np.uint64(24)) + cc[:nb]) & m48
        out[pos:pos + nb] = st.astype(np.float64) * (1.0 / (1 << 48))
        x = int(st[nb - 1])
        pos += nb
    return out


def drand48_first(seeds):
    """drand48() right after srand48(seed), vectorised over an integer array of seeds."""
    s = np.asarray(seeds, dtype=np.int64).astype(np.uint64)
    x = ((s & np.uint64(0xFFFFFFFF)) << np.uint64(16)) | np.uint64(0x330E)
    x_lo, x_hi = x & np.uint64(0xFFFFFF), x >> np.uint64(24)
    a_lo, a_hi = np.uint64(_A & 0xFFFFFF), np.uint64(_A >> 24)
    mid = (a_lo * x_hi + a_hi * x_lo) & np.uint64(0xFFFFFF)
    st = (a_lo * x_lo + (mid << np.uint64(24)) + np.uint64(_C)) & np.uint64(_MASK)
    return st.astype(np.float64) * (1.0 / (1 << 48))


def local_dim(global_dim, grid):
    """ceil(global/grid): matrix ctor, src/matrix/matrix.hpp:8-11."""
    return global_dim // grid + (1 if global_dim % grid else 0)


def _pad_len(global_dim, grid, coord, ldim):
    """padXlen / padYlen of structure.hpp:73-74,110-111."""
    if global_dim % grid != 0 and (ldim - 1) * grid + coord >= global_dim:
        return ldim - 1
    return ldim


def distribute_random(gx, gy, px, py, grid_x, grid_y, key):
    """rect::_distribute_random, src/matrix/structure.hpp:106-129.

    gx/gy: global columns/rows; (px,py): this rank's grid coordinate; one sequential drand48
    stream per rank, consumed column by column over the non-padded local block."""
    lx, ly = local_dim(gx, grid_x), local_dim(gy, grid_y)
    padx, pady = _pad_len(gx, grid_x, px, lx), _pad_len(gy, grid_y, py, ly)
    a = np.zeros((ly, lx), order="F")
    vals = drand48_stream(key, padx * pady)
    a[:pady, :padx] = vals.reshape((pady, padx), order="F")
    return a


def distribute_symmetric(gx, gy, px, py, grid_x, grid_y, diagonally_dominant=True):
    """rect::_distribute_symmetric, src/matrix/structure.hpp:69-103: element (row gy_, col gx_) is
    the first drand48 after srand48(max + globalDimY*min); +globalDimX on the global diagonal."""
    lx, ly = local_dim(gx, grid_x), local_dim(gy, grid_y)
    padx, pady = _pad_len(gx, grid_x, px, lx), _pad_len(gy, grid_y, py, ly)
    a = np.zeros((ly, lx), order="F")
    cols = px + grid_x * np.arange(padx, dtype=np.int64)
    rows = py + grid_y * np.arange(pady, dtype=np.int64)
    R, C = np.meshgrid(rows, cols, indexing="ij")
    seeds = np.maximum(R, C) + gy * np.minimum(R, C)
    blk = drand48_first(seeds)
    if diagonally_dominant:
        # reference adds only where global x == global y AND local i == j (structure.hpp:87)
        ii, jj = np.meshgrid(np.arange(pady), np.arange(padx), indexing="ij")
        blk = blk + np.where((R == C) & (ii == jj), float(gx), 0.0)
    a[:pady, :padx] = blk
    return a


# ----------------------------------------------------------------------------- engines
def syrk_upper_trans(a):
    """blas::engine::_syrk U/Trans alpha=1 beta=0 (src/blas/interface.hpp:82-97 <- cacqr.hpp:15):
    upper triangle of A^T A; the strict lower part is NOT written (returned as zeros here)."""
    return np.triu(a.T @ a)


def potrf_upper(g):
    """lapack::engine::_potrf U (src/lapack/interface.hpp:31-43): R upper with R^T R = G, reading
    only the upper triangle of G.  Unblocked right-looking restatement of LAPACK dpotrf."""
    n = g.shape[0]
    r = np.triu(np.array(g, dtype=np.float64, order="F"))
    for j in range(n):
        d = r[j, j]
        if not d > 0.0:
            raise np.linalg.LinAlgError("potrf: leading minor %d not positive definite" % (j + 1))
        d = np.sqrt(d)
        r[j, j] = d
        if j + 1 < n:
            r[j, j + 1:] /= d
            row = r[j, j + 1:]
            r[j + 1:, j + 1:] -= np.triu(np.outer(row, row))
    return r


def trtri_upper(r):
    """lapack::engine::_trtri U/NonUnit (src/lapack/interface.hpp:46-58): inverse of the upper
    triangle, column-by-column back substitution (LAPACK dtrti2 order)."""
    n = r.shape[0]
    t = np.triu(np.array(r, dtype=np.float64, order="F"))
    inv = np.zeros_like(t)
    for j in range(n):
        inv[j, j] = 1.0 / t[j, j]
        if j:
            # inv[0:j, j] = -inv[0:j,0:j] @ t[0:j, j] * inv[j,j]
            inv[:j, j] = -(inv[:j, :j] @ t[:j, j]) * inv[j, j]
    return inv


def trmm_right_upper(b, t):
    """_trmm Right/Upper/NoTrans/NonUnit alpha=1 (cacqr.hpp:24-25,184-187): B <- B * triu(T)."""
    return b @ np.triu(t)


# ----------------------------------------------------------------------------- CholeskyQR2 (1D)
def cacqr_sweep_1d(q_blocks):
    """qr::cacqr::sweep_1d, src/alg/qr/cacqr/cacqr.hpp:7-29, over P row blocks (one per rank):
    syrk -> allreduce (summed in rank order, like the oracle mini-MPI) -> potrf -> trtri -> trmm."""
    gram = syrk_upper_trans(q_blocks[0])
    for blk in q_blocks[1:]:
        gram = gram + syrk_upper_trans(blk)      # policy.h:82 MPI_Allreduce(SUM, world)
    r = potrf_upper(gram)                        # cacqr.hpp:20
    rinv = trtri_upper(r)                        # cacqr.hpp:21-22
    return [trmm_right_upper(blk, rinv) for blk in q_blocks], r   # cacqr.hpp:25


def cacqr_factor_1d(a_blocks, num_iter=2):
    """qr::cacqr::factor -> invoke_1d (cacqr.hpp:219-248,174-193): CholeskyQR (num_iter=1) or
    CholeskyQR2 (num_iter=2) of the row-partitioned matrix; returns (Q blocks, R upper n x n)."""
    q, r1 = cacqr_sweep_1d([np.array(b, order="F") for b in a_blocks])
    if num_iter > 1:
        q, r2 = cacqr_sweep_1d(q)
        r1 = trmm_right_upper(r2, r1)            # cacqr.hpp:184-187: R <- R2 * R1
    return q, np.triu(r1)


def cacqr_input_blocks(m, n, ranks):
    """bench/qr/cacqr.cpp:32-34 at c=1: matrix A(n, m, 1, P); distribute_random(0, rank, 1, P, key=rank)."""
    return [distribute_random(n, m, 0, r, 1, ranks, r) for r in range(ranks)]


def assemble_row_cyclic(blocks, m):
    """global row = rank + P*local row (element-cyclic deal, structure.hpp:74-98)."""
    p = len(blocks)
    out = np.zeros((m, blocks[0].shape[1]), order="F")
    for r, blk in enumerate(blocks):
        rows = np.arange(r, m, p)
        out[rows, :] = blk[:len(rows), :]
    return out


# ----------------------------------------------------------------------------- cholinv (P=1 schedule)
def cholinv_factor(a, complete_inv=True, split=1, bc_dim=None):
    """cholesky::cholinv::factor/invoke on one rank (src/alg/cholesky/cholinv/cholinv.hpp:6-47,87-165).

    Recursion on the upper factor: left block, R12 = R11^{-T} A12 (explicit inverse + trmm,
    :116-121), A22 -= R12^T R12 (:131-134), right block, then the inverse completion
    Rinv12 = -R11^{-1} R12 R22^{-1} (:147-156), skipped at the top level when !complete_inv.
    Base case (policy.h:196-201): potrf then trtri of the diagonal block.
    bc_dim = args.bcDimension (cholinv.hpp:15-18).  Returns (R, Rinv) upper triangular."""
    n = a.shape[0]
    work = np.triu(np.array(a, dtype=np.float64, order="F"))
    rinv = np.zeros_like(work)
    if bc_dim is None:
        bc_dim = n

    def rec(s, e, top):
        ld = e - s
        s1 = ld >> split
        if ld <= bc_dim or s1 < split:
            r = potrf_upper(work[s:e, s:e])
            work[s:e, s:e] = r
            rinv[s:e, s:e] = trtri_upper(r)
            return
        m = s + s1
        rec(s, m, False)
        work[s:m, m:e] = np.triu(rinv[s:m, s:m]).T @ work[s:m, m:e]
        work[m:e, m:e] -= np.triu(work[s:m, m:e].T @ work[s:m, m:e])
        rec(m, e, False)
        if not (not complete_inv and top):
            x = np.triu(rinv[s:m, s:m]) @ work[s:m, m:e]
            rinv[s:m, m:e] = -(x @ np.triu(rinv[m:e, m:e]))

    rec(0, n, True)
    return work, rinv


def cholinv_bc_dimension(n, c, d, bc_mult_dim):
    """Base-case size formula, cholinv.hpp:15-18."""
    ldim = local_dim(n, d)
    bc = c * d
    if bc_mult_dim < 0:
        bc *= 2 ** (-bc_mult_dim)
    else:
        bc //= 2 ** bc_mult_dim
    bc = min(ldim, max(1, bc))
    return d * (ldim // bc)


def assemble_cyclic_2d(blocks, n_rows, n_cols, d):
    """blocks[(x, y)] -> global matrix; global (row, col) = (y + d*row_loc, x + d*col_loc)."""
    out = np.zeros((n_rows, n_cols), order="F")
    for (x, y), blk in blocks.items():
        rows, cols = np.arange(y, n_rows, d), np.arange(x, n_cols, d)
        out[np.ix_(rows, cols)] = blk[:len(rows), :len(cols)]
    return out


def symmetric_global(n):
    """Global input of bench/cholesky/cholinv.cpp:39-40 (grid-independent)."""
    return distribute_symmetric(n, n, 0, 0, 1, 1, True)


# ----------------------------------------------------------------------------- SUMMA
def summa_gemm(a, b, c=None, alpha=1.0, beta=0.0, ta=False, tb=False):
    """matmult::summa::invoke gemm overload (src/alg/matmult/summa/summa.hpp:7-44) seen globally:
    C = alpha*op(A)*op(B) + beta*C (beta applied after the depth allreduce, :32-35)."""
    prod = alpha * ((a.T if ta else a) @ (b.T if tb else b))
    return prod if (c is None or beta == 0.0) else beta * c + prod


# ----------------------------------------------------------------------------- reference metrics
def qr_orthogonality(q):
    """qr::validate::orthogonality (test/qr/validate.hpp:7-32 + util.hpp:27-53): ||Q^T Q - I||_F / sqrt(n^2)."""
    n = q.shape[1]
    return np.linalg.norm(q.T @ q - np.eye(n)) / np.sqrt(n * n)


def qr_residual(a, q, r):
    """qr::validate::residual (test/qr/validate.hpp:37-52): ||QR - A||_F / ||A||_F."""
    return np.linalg.norm(q @ np.triu(r) - a) / np.linalg.norm(a)


def cholesky_residual(a, r):
    """cholesky::validate::residual, dir 'U' (test/cholesky/validate.hpp:31-46): upper triangle only."""
    ru = np.triu(r)
    return np.linalg.norm(np.triu(ru.T @ ru - a)) / np.linalg.norm(np.triu(a))


# ----------------------------------------------------------------------------- flop counts
def flops_cacqr2(m, n):
    return 4.0 * m * n * n + (5.0 / 3.0) * n ** 3


def flops_cholesky(n):
    return n ** 3 / 3.0


def flops_gemm(m, n, k):
    return 2.0 * m * n * k
