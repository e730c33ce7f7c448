"""ctypes binding of include/capital_b200.h.  Fails loudly when the CUDA library is missing: there is
no CPU or PyTorch fallback behind these calls."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcapital_b200.so")

_i64, _int, _dbl, _ptr, _chr = C.c_int64, C.c_int, C.c_double, C.c_void_p, C.c_char

_SIGS = {
    "cap_create": [C.POINTER(_ptr), _int], "cap_destroy": [_ptr], "cap_set_stream": [_ptr, _ptr],
    "cap_synchronize": [_ptr], "cap_set_engine": [_ptr, _int], "cap_set_host_chunk_bytes": [_ptr, C.c_size_t], "cap_use_own_stream": [_ptr], "cap_release_workspace": [_ptr], "cap_set_cacqr_complete_inv": [_ptr, _int], "cap_set_peer_exchange": [_ptr, _int], "cap_set_update_grid": [_ptr, _int], "cap_get_info": [_ptr, C.POINTER(_int)], "cap_reset_info": [_ptr],
    "cap_malloc": [C.POINTER(_ptr), C.c_size_t], "cap_free": [_ptr], "cap_memset_async": [_ptr, _ptr, _int, C.c_size_t],
    "cap_h2d_async": [_ptr, _ptr, _ptr, C.c_size_t], "cap_d2h_async": [_ptr, _ptr, _ptr, C.c_size_t],
    "cap_host_alloc": [C.POINTER(_ptr), C.c_size_t], "cap_host_free": [_ptr],
    "cap_dgemm": [_ptr, _chr, _chr, _i64, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64, _dbl, _ptr, _i64],
    "cap_dsyrk": [_ptr, _chr, _chr, _i64, _i64, _dbl, _ptr, _i64, _dbl, _ptr, _i64],
    "cap_dtrmm": [_ptr, _chr, _chr, _chr, _chr, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64],
    "cap_dtrmm_oop": [_ptr, _chr, _chr, _chr, _chr, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64, _ptr, _i64],
    "cap_dtrsm": [_ptr, _chr, _chr, _chr, _chr, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64],
    "cap_set_block_sizes": [_ptr, _int, _int], "cap_set_lookahead": [_ptr, _int, _int], "cap_set_potrf_schedule": [_ptr, _int],
    "cap_potrf_1d": [_ptr, _ptr, _i64, _int, _ptr, _i64],
    "cap_potrf_1d_layout": [_i64, _int, _int, _int, C.POINTER(_i64), C.POINTER(_int)],
    "cap_potrf_1d_block": [_i64, _int, C.POINTER(_int), C.POINTER(_int)],
    "cap_dpotrf": [_ptr, _chr, _int, _ptr, _int], "cap_dtrtri": [_ptr, _chr, _chr, _int, _ptr, _int],
    "cap_dcholinv": [_ptr, _int, _ptr, _i64, _ptr, _i64, _int],
    "cap_cqr_gram": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64], "cap_cqr_gram_noalloc": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64],
    "cap_cqr_apply": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64],
    "cap_copy2d": [_ptr, _ptr, _i64, _ptr, _i64, _i64, _i64, _int],
    "cap_pack_upper": [_ptr, _ptr, _ptr, _i64, _int], "cap_unpack_upper": [_ptr, _ptr, _i64, _ptr, _int, _int],
    "cap_axpby": [_ptr, _ptr, _ptr, _i64, _dbl, _dbl],
    "cap_axpby2d": [_ptr, _ptr, _i64, _ptr, _i64, _i64, _i64, _dbl, _dbl], "cap_fill_diagonal": [_ptr, _ptr, _i64, _i64, _dbl],
    "cap_fill_random": [_ptr, _ptr, _i64, _i64, _i64, C.c_longlong],
    "cap_fill_symmetric": [_ptr, _ptr, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _int],
    "cap_comm_unique_id": [C.c_char_p], "cap_comm_init_rank": [C.POINTER(_ptr), C.c_char_p, _int, _int],
    "cap_comm_split": [_ptr, _int, _int, C.POINTER(_ptr)], "cap_comm_destroy": [_ptr],
    "cap_comm_rank": [_ptr, C.POINTER(_int)], "cap_comm_size": [_ptr, C.POINTER(_int)],
    "cap_allreduce_sum": [_ptr, _ptr, _ptr, _i64], "cap_reduce_sum": [_ptr, _ptr, _ptr, _i64, _int],
    "cap_bcast": [_ptr, _ptr, _ptr, _i64, _int], "cap_allgather": [_ptr, _ptr, _ptr, _ptr, _i64],
    "cap_sendrecv": [_ptr, _ptr, _ptr, _ptr, _i64, _int],
    "cap_grid_create": [_ptr, _int, C.POINTER(_ptr)], "cap_grid_create2": [_ptr, _int, _int, _int, C.POINTER(_ptr)], "cap_grid_set_chunks": [_ptr, _int], "cap_grid_destroy": [_ptr],
    "cap_grid_info": [_ptr] + [C.POINTER(_int)] * 5, "cap_grid_transpose": [_ptr, _ptr, _ptr, _ptr, _i64],
    "cap_summa_gemm": [_ptr, _ptr, _chr, _chr, _i64, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64, _dbl, _ptr, _i64],
    "cap_summa_trmm": [_ptr, _ptr, _chr, _chr, _chr, _chr, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64],
    "cap_summa_syrk": [_ptr, _ptr, _chr, _chr, _i64, _i64, _dbl, _ptr, _i64, _ptr, _i64, _dbl, _ptr, _i64],
    "cap_cholinv_factor": [_ptr, _ptr, _i64, _ptr, _i64, _ptr, _i64, _int, _int, _i64],
    "cap_cacqr_factor_1d": [_ptr, _ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _int],
    "cap_cacqr_factor_1d_host": [_ptr, _ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _int],
    "cap_residual_begin": [_ptr],
    "cap_residual_accumulate": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _int, _int, _int, _int, _dbl],
    "cap_residual_finish": [_ptr, _ptr, _ptr],
    "cap_validate_cacqr_1d": [_ptr, _ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _ptr],
    "cap_validate_cacqr_3d": [_ptr, _ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _ptr],
    "cap_validate_cholinv": [_ptr, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr],
    "cap_cacqr_factor_3d": [_ptr, _ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _int, _int, _i64],
}
EXPORTS = sorted(list(_SIGS) + ["cap_last_error", "cap_get_stream", "cap_kernel_launches", "cap_debug_stair_decode"])

_lib = None


class CapitalError(RuntimeError):
    pass


def load():
    """dlopen the C-ABI library (torch first, so that its bundled libnccl.so.2 is the one resolved)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CapitalError("capital_b200: %s is missing -- run `python -m capital_b200.build` (nvcc, sm_100a). "
                           "There is no CPU fallback." % LIB_PATH)
    import torch  # noqa: F401  (loads libnccl.so.2 / libcudart)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, args in _SIGS.items():
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = args, _int
    lib.cap_last_error.restype = C.c_char_p
    lib.cap_get_stream.argtypes, lib.cap_get_stream.restype = [_ptr], _ptr
    lib.cap_kernel_launches.argtypes, lib.cap_kernel_launches.restype = [_ptr], C.c_longlong
    lib.cap_debug_stair_decode.argtypes = [_int, _int, _int, _int, _int, _int, _i64, C.POINTER(_int), C.POINTER(_int), C.POINTER(_int)]
    lib.cap_debug_stair_decode.restype = _i64
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        raise CapitalError("%s failed (status %d): %s" % (what, rc, load().cap_last_error().decode()))
