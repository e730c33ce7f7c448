"""Python host mirror of the reference's engine interface over the C ABI.

Names and argument meaning follow blas::engine::_gemm/_trmm/_syrk (reference src/blas/interface.h:57-66)
and lapack::engine::_potrf/_trtri (src/lapack/interface.h:49-53); matrices are torch CUDA float64
tensors in COLUMN-MAJOR layout (stride(0) == 1), the reference's only layout on this path.
torch is plumbing here (device memory + streams); all arithmetic is in libcapital_b200.so."""
import ctypes as C

import torch

from . import _lib


def colmajor(rows, cols, device="cuda", fill=None):
    """rows x cols float64 tensor with Fortran strides (1, rows)."""
    base = torch.empty((cols, rows), dtype=torch.float64, device=device)
    if fill is not None:
        base.fill_(fill)
    return base.t()


def to_colmajor(t):
    """Copy any 2-D tensor/ndarray into a column-major CUDA tensor."""
    t = torch.as_tensor(t, dtype=torch.float64)
    out = colmajor(t.shape[0], t.shape[1], device="cuda")
    out.copy_(t)
    return out


def _ld(t):
    assert t.dtype == torch.float64 and t.is_cuda and t.dim() == 2, "need a 2-D CUDA float64 tensor"
    assert t.shape[0] <= 1 or t.stride(0) == 1, "matrix must be column-major (stride(0) == 1)"
    return max(1, t.stride(1)) if t.shape[1] > 1 else max(1, t.shape[0])


def _c(ch):
    return C.c_char(ch.encode())


class Engine:
    """One handle = one device + stream + library-owned workspace (the reference's engines are static
    classes; the handle is what a device back end needs in addition)."""

    def __init__(self, device=0):
        self.lib = _lib.load()
        self.h = C.c_void_p()
        _lib.check(self.lib.cap_create(C.byref(self.h), int(device)), "cap_create")
        self.device = int(device)
        self.use_torch_stream()

    def use_torch_stream(self):
        s = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.cap_set_stream(self.h, C.c_void_p(s)), "cap_set_stream")

    def set_engine(self, engine):
        """0 = auto (TMA engine when alignment allows), 1 = cp.async engine only."""
        _lib.check(self.lib.cap_set_engine(self.h, int(engine)), "cap_set_engine")

    def set_host_chunk_bytes(self, nbytes):
        _lib.check(self.lib.cap_set_host_chunk_bytes(self.h, int(nbytes)), "cap_set_host_chunk_bytes")

    def release_workspace(self):
        """Synchronise and free the library-owned scratch (grown on demand, otherwise kept until close())."""
        _lib.check(self.lib.cap_release_workspace(self.h), "cap_release_workspace")

    def close(self):
        if self.h:
            self.lib.cap_destroy(self.h)
            self.h = C.c_void_p()

    def launches(self):
        return int(self.lib.cap_kernel_launches(self.h))

    def info(self):
        v = C.c_int(0)
        _lib.check(self.lib.cap_get_info(self.h, C.byref(v)), "cap_get_info")
        return v.value

    def reset_info(self):
        _lib.check(self.lib.cap_reset_info(self.h), "cap_reset_info")

    # ---- blas::engine -------------------------------------------------------------------------
    def _gemm(self, A, B, Cm, m, n, k, transa="N", transb="N", alpha=1.0, beta=0.0):
        _lib.check(self.lib.cap_dgemm(self.h, _c(transa), _c(transb), m, n, k, alpha, A.data_ptr(), _ld(A),
                                      B.data_ptr(), _ld(B), beta, Cm.data_ptr(), _ld(Cm)), "cap_dgemm")

    def _syrk(self, A, Cm, n, k, uplo="U", trans="T", alpha=1.0, beta=0.0):
        _lib.check(self.lib.cap_dsyrk(self.h, _c(uplo), _c(trans), n, k, alpha, A.data_ptr(), _ld(A), beta,
                                      Cm.data_ptr(), _ld(Cm)), "cap_dsyrk")

    def _trmm(self, T, B, m, n, side="R", uplo="U", trans="N", diag="N", alpha=1.0):
        _lib.check(self.lib.cap_dtrmm(self.h, _c(side), _c(uplo), _c(trans), _c(diag), m, n, alpha, T.data_ptr(), _ld(T),
                                      B.data_ptr(), _ld(B)), "cap_dtrmm")

    def trmm_oop(self, T, B, Cm, m, n, side="R", uplo="U", trans="N", diag="N", alpha=1.0):
        _lib.check(self.lib.cap_dtrmm_oop(self.h, _c(side), _c(uplo), _c(trans), _c(diag), m, n, alpha, T.data_ptr(), _ld(T),
                                          B.data_ptr(), _ld(B), Cm.data_ptr(), _ld(Cm)), "cap_dtrmm_oop")

    def trsm(self, T, B, m, n, side="L", uplo="U", trans="N", diag="N", alpha=1.0):
        """op(T) X = alpha B (side L) or X op(T) = alpha B (side R), in place on B (cap_dtrsm)."""
        _lib.check(self.lib.cap_dtrsm(self.h, _c(side), _c(uplo), _c(trans), _c(diag), m, n, alpha, T.data_ptr(), _ld(T),
                                      B.data_ptr(), _ld(B)), "cap_dtrsm")

    def set_block_sizes(self, potrf_nb=0, trsm_nb=512):
        _lib.check(self.lib.cap_set_block_sizes(self.h, int(potrf_nb), int(trsm_nb)), "cap_set_block_sizes")

    def set_lookahead(self, reserve_sms=16, pipeline=-1):
        _lib.check(self.lib.cap_set_lookahead(self.h, int(reserve_sms), int(pipeline)), "cap_set_lookahead")

    def set_cacqr_complete_inv(self, complete_inv=True):
        """3D CholeskyQR: True = completed inverse, False = cacqr::solve (cacqr.hpp:46-73)."""
        _lib.check(self.lib.cap_set_cacqr_complete_inv(self.h, int(bool(complete_inv))), "cap_set_cacqr_complete_inv")

    def set_peer_exchange(self, mode=-1):
        """Panel exchange of potrf_1d: -1 = peer copy engines when CUDA IPC works, 0 = NCCL only, 1 = peer copy engines required."""
        _lib.check(self.lib.cap_set_peer_exchange(self.h, int(mode)), "cap_set_peer_exchange")

    def set_update_grid(self, ctas_per_sm=0):
        """Trailing update beside a chain: 0 = persistent SM-capped grid, k > 0 = k x #SMs CTAs without a cap."""
        _lib.check(self.lib.cap_set_update_grid(self.h, int(ctas_per_sm)), "cap_set_update_grid")

    def set_potrf_schedule(self, schedule):
        """0 = right-looking with look-ahead (cap_potrf_1d on one rank), 1 = recursive potrf / trsm / syrk."""
        _lib.check(self.lib.cap_set_potrf_schedule(self.h, int(schedule)), "cap_set_potrf_schedule")

    def potrf_1d(self, A_loc, N, nb, comm=None):
        """Look-ahead Cholesky of the global N x N matrix whose block columns (capital_b200.dist.potrf_1d_blocks) this rank
        holds side by side in A_loc (N x ncols_loc, column-major); comm = the raw cap_comm_t or None."""
        _lib.check(self.lib.cap_potrf_1d(self.h, comm, N, nb, A_loc.data_ptr(), _ld(A_loc)), "cap_potrf_1d")

    # ---- lapack::engine -----------------------------------------------------------------------
    def _potrf(self, A, n, uplo="U"):
        _lib.check(self.lib.cap_dpotrf(self.h, _c(uplo), n, A.data_ptr(), _ld(A)), "cap_dpotrf")

    def _trtri(self, A, n, uplo="U", diag="N"):
        _lib.check(self.lib.cap_dtrtri(self.h, _c(uplo), _c(diag), n, A.data_ptr(), _ld(A)), "cap_dtrtri")

    def cholinv(self, A, Rinv, n, complete_inv=True):
        _lib.check(self.lib.cap_dcholinv(self.h, n, A.data_ptr(), _ld(A), Rinv.data_ptr(), _ld(Rinv), int(bool(complete_inv))), "cap_dcholinv")

    # ---- CholeskyQR -----------------------------------------------------------------------------
    def cqr_gram(self, Q, G):
        m, n = Q.shape
        _lib.check(self.lib.cap_cqr_gram(self.h, m, n, Q.data_ptr(), _ld(Q), G.data_ptr(), _ld(G)), "cap_cqr_gram")

    def cqr_apply(self, Rinv, Qin, Qout):
        m, n = Qin.shape
        _lib.check(self.lib.cap_cqr_apply(self.h, m, n, Rinv.data_ptr(), _ld(Rinv), Qin.data_ptr(), _ld(Qin),
                                          Qout.data_ptr(), _ld(Qout)), "cap_cqr_apply")

    def cacqr_factor_1d(self, A, Q, R, num_iter=2, comm=None):
        m, n = A.shape
        _lib.check(self.lib.cap_cacqr_factor_1d(self.h, comm, m, n, A.data_ptr(), _ld(A), Q.data_ptr(), _ld(Q),
                                                R.data_ptr(), _ld(R), num_iter), "cap_cacqr_factor_1d")

    def cacqr_factor_1d_host(self, A_host, Q_host, R_host, m, n, num_iter=2, comm=None):
        """A_host/Q_host: pinned CPU tensors holding the m x n block column-major (shape (n, m) contiguous)."""
        _lib.check(self.lib.cap_cacqr_factor_1d_host(self.h, comm, m, n, A_host.data_ptr(), m, Q_host.data_ptr(), m,
                                                     R_host.data_ptr(), n, num_iter), "cap_cacqr_factor_1d_host")

    # ---- process grid: summa / cholinv (all dimensions local) --------------------------------------
    def summa_gemm(self, grid, A, B, Cm, m, n, k, transa="N", transb="N", alpha=1.0, beta=0.0):
        _lib.check(self.lib.cap_summa_gemm(self.h, grid.ptr, _c(transa), _c(transb), m, n, k, alpha, A.data_ptr(), _ld(A),
                                           B.data_ptr(), _ld(B), beta, Cm.data_ptr(), _ld(Cm)), "cap_summa_gemm")

    def summa_trmm(self, grid, T, B, m, n, side="L", uplo="U", trans="N", diag="N", alpha=1.0):
        _lib.check(self.lib.cap_summa_trmm(self.h, grid.ptr, _c(side), _c(uplo), _c(trans), _c(diag), m, n, alpha,
                                           T.data_ptr(), _ld(T), B.data_ptr(), _ld(B)), "cap_summa_trmm")

    def summa_syrk(self, grid, A, Cm, n, k, uplo="U", trans="T", alpha=1.0, beta=0.0, At=None):
        _lib.check(self.lib.cap_summa_syrk(self.h, grid.ptr, _c(uplo), _c(trans), n, k, alpha, A.data_ptr(), _ld(A),
                                           At.data_ptr() if At is not None else None, _ld(At) if At is not None else 0,
                                           beta, Cm.data_ptr(), _ld(Cm)), "cap_summa_syrk")

    def grid_transpose(self, grid, src, dst):
        _lib.check(self.lib.cap_grid_transpose(self.h, grid.ptr, src.data_ptr(), dst.data_ptr(), src.numel()), "cap_grid_transpose")

    def cholinv_factor(self, grid, A, Rinv, complete_inv=True, split=1, bc_global=2048):
        n = A.shape[0]
        _lib.check(self.lib.cap_cholinv_factor(self.h, grid.ptr, n, A.data_ptr(), _ld(A), Rinv.data_ptr(), _ld(Rinv),
                                               int(bool(complete_inv)), int(split), int(bc_global)), "cap_cholinv_factor")

    def cacqr_factor_3d(self, grid, A, Q, R, num_iter=2, Rinv=None, split=1, bc_global=2048):
        """qr::cacqr::factor on a process grid (cacqr.hpp:219-248, 3D branch): local m_loc x n_loc blocks."""
        m_loc, n_loc = A.shape
        _lib.check(self.lib.cap_cacqr_factor_3d(self.h, grid.ptr, m_loc, n_loc, A.data_ptr(), _ld(A), Q.data_ptr(), _ld(Q),
                                                R.data_ptr(), _ld(R), Rinv.data_ptr() if Rinv is not None else None,
                                                _ld(Rinv) if Rinv is not None else 0, int(num_iter), int(split), int(bc_global)),
                   "cap_cacqr_factor_3d")

    # ---- validators (test/qr/validate.hpp, test/cholesky/validate.hpp), computed on the device ------------
    def validate_cacqr_1d(self, A, Q, R, comm=None):
        """(||Q^T Q - I||_F / n, ||Q R - A||_F / ||A||_F) for the 1D algorithm; comm = the raw cap_comm_t or None."""
        m, n = A.shape
        orth, resid = C.c_double(), C.c_double()
        _lib.check(self.lib.cap_validate_cacqr_1d(self.h, comm, m, n, A.data_ptr(), _ld(A), Q.data_ptr(), _ld(Q), R.data_ptr(), _ld(R),
                                                  C.byref(orth), C.byref(resid)), "cap_validate_cacqr_1d")
        return orth.value, resid.value

    def validate_cacqr_3d(self, grid, A, Q, R):
        m, n = A.shape
        orth, resid = C.c_double(), C.c_double()
        _lib.check(self.lib.cap_validate_cacqr_3d(self.h, grid.ptr, m, n, A.data_ptr(), _ld(A), Q.data_ptr(), _ld(Q), R.data_ptr(), _ld(R),
                                                  C.byref(orth), C.byref(resid)), "cap_validate_cacqr_3d")
        return orth.value, resid.value

    def validate_cholinv(self, grid, A, R):
        """||R^T R - A||_F / ||A||_F over the upper triangle; A = the original local block."""
        resid = C.c_double()
        _lib.check(self.lib.cap_validate_cholinv(self.h, grid.ptr, A.shape[0], A.data_ptr(), _ld(A), R.data_ptr(), _ld(R), C.byref(resid)),
                   "cap_validate_cholinv")
        return resid.value

    # ---- serialize / generators -----------------------------------------------------------------
    def copy2d(self, dst, src, m, n, mode=0):
        _lib.check(self.lib.cap_copy2d(self.h, dst.data_ptr(), _ld(dst), src.data_ptr(), _ld(src), m, n, mode), "cap_copy2d")

    def pack_upper(self, packed, full, n):
        _lib.check(self.lib.cap_pack_upper(self.h, packed.data_ptr(), full.data_ptr(), _ld(full), n), "cap_pack_upper")

    def unpack_upper(self, full, packed, n, zero_lower=True):
        _lib.check(self.lib.cap_unpack_upper(self.h, full.data_ptr(), _ld(full), packed.data_ptr(), n, int(zero_lower)), "cap_unpack_upper")

    def axpby(self, y, x, alpha, beta):
        _lib.check(self.lib.cap_axpby(self.h, y.data_ptr(), x.data_ptr(), y.numel(), alpha, beta), "cap_axpby")

    def fill_random(self, dst, key):
        """matrix::distribute_random (reference src/matrix/structure.hpp:106-129) for this rank's local block."""
        _lib.check(self.lib.cap_fill_random(self.h, dst.data_ptr(), _ld(dst), dst.shape[0], dst.shape[1], int(key)), "cap_fill_random")

    def fill_symmetric(self, dst, px, py, grid_x, grid_y, global_dim, diagonally_dominant=True):
        """matrix::distribute_symmetric (src/matrix/structure.hpp:69-103)."""
        _lib.check(self.lib.cap_fill_symmetric(self.h, dst.data_ptr(), _ld(dst), dst.shape[0], dst.shape[1], px, py, grid_x, grid_y,
                                               global_dim, int(diagonally_dominant)), "cap_fill_symmetric")
