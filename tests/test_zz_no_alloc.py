"""GPU: cap_cqr_gram_noalloc -- the allocation-free variant of cap_cqr_gram (split-K partials in a static device array,
include/capital_b200.h) -- against the existing entry point, through the C ABI, on two calls in a row.  No counterpart
in the reference.  Collected last on purpose;
passes alone in a fresh process and after the other tests (it uses its own handle)."""
import ctypes as C

import pytest

from conftest import TOL_FACTOR

pytestmark = pytest.mark.gpu


def _gram(lib, fn, h, Q, G):
    from capital_b200 import _lib
    m, n = Q.shape
    _lib.check(fn(h, m, n, Q.data_ptr(), max(1, Q.stride(1)), G.data_ptr(), max(1, G.stride(1))), "gram")


@pytest.mark.parametrize("m,n", [(65536, 256), (40000, 512), (3001, 130), (100, 64)])
def test_gram_noalloc_matches_gram_twice(m, n):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from capital_b200 import _lib
    from capital_b200.engine import Engine, colmajor
    eng = Engine(0)                      # own handle: the test does not depend on what ran before it
    lib = eng.lib
    try:
        for call in range(2):            # the second call finds the first call's partials in the shared static array
            Q = colmajor(m, n)
            eng.fill_random(Q, 17 + call)
            if call == 1:
                Q.mul_(3.0).sub_(1.0)
            want = colmajor(n, n, fill=0.0); got = colmajor(n, n, fill=0.0)
            _gram(lib, lib.cap_cqr_gram, eng.h, Q, want)
            torch.cuda.synchronize()
            _gram(lib, lib.cap_cqr_gram_noalloc, eng.h, Q, got)           # (also loads the kernels: lazy module loading is not an allocation of ours)
            torch.cuda.synchronize()
            got.zero_()
            free_before = torch.cuda.mem_get_info()[0]
            _gram(lib, lib.cap_cqr_gram_noalloc, eng.h, Q, got)
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info()[0] == free_before, "the allocation-free entry point changed the free device memory"
            a, b = torch.triu(got), torch.triu(want)
            rel = float((a - b).norm() / b.norm())
            assert rel <= TOL_FACTOR, (call, rel)
            assert torch.equal(torch.tril(got, -1), torch.zeros_like(got).tril(-1)), "strict lower triangle of G must not be written"
            ref = torch.triu(Q.t() @ Q)                       # independent check of both entry points
            assert float((a - ref).norm() / ref.norm()) <= TOL_FACTOR
    finally:
        eng.close()


def test_gram_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from capital_b200.engine import Engine, colmajor
    eng = Engine(0)
    try:
        n = 513
        Q = colmajor(1024, n, fill=1.0); G = colmajor(n, n, fill=0.0)
        rc = eng.lib.cap_cqr_gram_noalloc(eng.h, 1024, n, Q.data_ptr(), 1024, G.data_ptr(), n)
        assert rc == 1, rc                                    # CAP_ERR_ARG
        eng.lib.cap_last_error.restype = C.c_char_p
        assert b"CAP_GRAM_NOALLOC_MAX_N" in eng.lib.cap_last_error()
        torch.cuda.synchronize()
        assert float(G.abs().sum()) == 0.0                    # nothing was launched
    finally:
        eng.close()
