"""GPU parity of the triangular-solve building block and the N^3/3 Cholesky built on it (cap_dtrsm, cap_dpotrf,
batched-leaf cap_dtrtri), through the C ABI, against numpy FP64 (LAPACK) on seeded well-conditioned inputs.
The reference has no _trsm engine call (src/blas/interface.h:57-66); _potrf / _trtri are lapack::engine
(src/lapack/interface.hpp:31-58).  Tolerances: BASELINE.json (relative factor difference <= 1e-11,
||A - R^T R|| / ||A|| <= 1e-13 * n)."""
import numpy as np
import pytest
import torch

import capital_oracle as orc
from conftest import TOL_FACTOR, TOL_RESIDUAL_PER_N, relerr

pytestmark = pytest.mark.gpu


def dev(a):
    from capital_b200.engine import to_colmajor
    return to_colmajor(np.asarray(a))


def host(t):
    torch.cuda.synchronize()
    return t.cpu().numpy()


def triangular(rng, n, uplo):
    """Well-conditioned triangle + garbage in the other half (must never be read)."""
    full = rng.random((n, n)) - 0.5
    tri = (np.triu(full) if uplo == "U" else np.tril(full)) + n * 0.25 * np.eye(n)
    junk = np.tril(full, -1) if uplo == "U" else np.triu(full, 1)
    return tri, tri + 1e3 * junk


@pytest.fixture
def blocks(engine):
    """Restore the default blocking after a test that shrinks it."""
    yield engine
    engine.set_block_sizes(0, 512); engine.set_potrf_schedule(0); engine.set_lookahead(16, -1)


@pytest.mark.parametrize("side", ["L", "R"])
@pytest.mark.parametrize("uplo", ["U", "L"])
@pytest.mark.parametrize("trans", ["N", "T"])
@pytest.mark.parametrize("nt,other,nb", [(1, 1, 512), (100, 37, 512), (128, 5, 128), (700, 300, 256), (1300, 260, 512), (641, 1030, 128)])
def test_trsm_vs_numpy(blocks, side, uplo, trans, nt, other, nb):
    eng = blocks
    eng.set_block_sizes(1024, nb)
    rng = np.random.default_rng(nt * 7 + other)
    tri, stored = triangular(rng, nt, uplo)
    b = rng.random((nt, other) if side == "L" else (other, nt)) - 0.5
    alpha = -1.5
    op = tri.T if trans == "T" else tri
    ref = np.linalg.solve(op, alpha * b) if side == "L" else np.linalg.solve(op.T, alpha * b.T).T
    T, B = dev(stored), dev(b)
    eng.trsm(T, B, b.shape[0], b.shape[1], side, uplo, trans, "N", alpha)
    assert relerr(host(B), ref) <= TOL_FACTOR
    assert np.array_equal(host(T), stored)                                   # the operand is not modified


def test_trsm_on_views_and_alpha_zero(engine):
    """Leading dimensions larger than the block (sub-matrix views), alpha == 0, empty right-hand side."""
    from capital_b200.engine import colmajor
    rng = np.random.default_rng(11)
    n, k = 300, 70
    tri, stored = triangular(rng, n, "U")
    big_t = rng.random((n + 13, n + 5)); big_t[3:3 + n, 2:2 + n] = stored
    big_b = rng.random((n + 9, k + 4)); b = big_b[5:5 + n, 1:1 + k].copy()
    T, B = dev(big_t), dev(big_b)
    engine.trsm(T[3:3 + n, 2:2 + n], B[5:5 + n, 1:1 + k], n, k, "L", "U", "T", "N", 1.0)
    out = host(B)
    assert relerr(out[5:5 + n, 1:1 + k], np.linalg.solve(tri.T, b)) <= TOL_FACTOR
    mask = np.ones_like(big_b, dtype=bool); mask[5:5 + n, 1:1 + k] = False
    assert np.array_equal(out[mask], big_b[mask])                            # nothing outside the view is touched
    Z = dev(rng.random((n, k)))
    engine.trsm(T[3:3 + n, 2:2 + n], Z, n, k, "L", "U", "N", "N", 0.0)
    assert not host(Z).any()
    engine.trsm(T[3:3 + n, 2:2 + n], colmajor(n, 1), n, 0, "L", "U", "N", "N", 1.0)   # n == 0: no-op


@pytest.mark.parametrize("n,nb", [(700, 128), (1000, 256), (1025, 1024), (2500, 1024), (3000, 512)])
def test_potrf_blocked_vs_numpy(blocks, n, nb):
    """cap_dpotrf on more than one diagonal block: recursive potrf / trsm / triangular syrk."""
    from capital_b200.engine import colmajor
    eng = blocks
    eng.set_block_sizes(nb, 512); eng.set_potrf_schedule(1)
    A = colmajor(n, n); eng.fill_symmetric(A, 0, 0, 1, 1, n, True)
    a_np = orc.symmetric_global(n)
    eng.reset_info(); eng._potrf(A, n, "U")
    out = host(A)
    R = np.linalg.cholesky(a_np).T
    assert eng.info() == 0
    assert relerr(np.triu(out), R) <= TOL_FACTOR
    assert orc.cholesky_residual(a_np, np.triu(out)) <= TOL_RESIDUAL_PER_N * n
    assert np.array_equal(np.tril(out, -1), np.tril(a_np, -1))               # strict lower untouched, like dpotrf


def test_potrf_blocked_random_spd_and_info(blocks):
    """A less structured SPD matrix (Gram of a random panel), then a non-positive pivot deep inside the second block."""
    eng = blocks
    eng.set_block_sizes(256, 512); eng.set_potrf_schedule(1)
    n = 900
    rng = np.random.default_rng(3)
    p = rng.random((2 * n, n)) - 0.5
    s = p.T @ p + 0.5 * np.eye(n)
    A = dev(s); eng.reset_info(); eng._potrf(A, n, "U")
    assert eng.info() == 0
    assert relerr(np.triu(host(A)), np.linalg.cholesky(s).T) <= 1e-10        # condition number ~1e2 here
    bad = s.copy(); bad[600, 600] = -1.0
    A = dev(bad); eng.reset_info(); eng._potrf(A, n, "U")
    assert eng.info() == 601                                                 # LAPACK: order of the first non-positive minor
    eng.reset_info()


@pytest.mark.parametrize("uplo", ["U", "L"])
@pytest.mark.parametrize("n", [129, 640, 1500])
def test_trtri_batched_leaves_vs_numpy(engine, uplo, n):
    rng = np.random.default_rng(n)
    tri, stored = triangular(rng, n, uplo)
    T = dev(stored); engine._trtri(T, n, uplo, "N")
    out = host(T)
    keep = np.triu if uplo == "U" else np.tril
    assert relerr(keep(out), np.linalg.inv(tri)) <= TOL_FACTOR
    other = np.tril(out, -1) if uplo == "U" else np.triu(out, 1)
    assert np.array_equal(other, np.tril(stored, -1) if uplo == "U" else np.triu(stored, 1))


@pytest.mark.parametrize("pipeline", [-1, 0, 1])
@pytest.mark.parametrize("n,nb,reserve", [(100, 128, 16), (640, 128, 16), (1000, 256, 0), (1300, 256, 16), (2500, 512, 16), (4096, 1024, 8), (1536, 128, 16)])
def test_potrf_lookahead_one_rank_vs_numpy(blocks, n, nb, reserve, pipeline):
    """cap_potrf_1d on one rank = cap_dpotrf with schedule 0: right-looking, panel(j+1) beside the trailing update of j."""
    from capital_b200.engine import colmajor
    eng = blocks
    eng.set_block_sizes(nb, 512); eng.set_lookahead(reserve, pipeline); eng.set_potrf_schedule(0)
    try:
        A = colmajor(n, n); eng.fill_symmetric(A, 0, 0, 1, 1, n, True)
        a_np = orc.symmetric_global(n)
        R = np.linalg.cholesky(a_np).T
        for _ in range(3):                                                   # repeated: the double-buffered panels and events are reused
            eng.fill_symmetric(A, 0, 0, 1, 1, n, True)
            eng.reset_info(); eng._potrf(A, n, "U")
            out = host(A)
            assert eng.info() == 0
            assert relerr(np.triu(out), R) <= TOL_FACTOR
            assert np.array_equal(np.tril(out, -1), np.tril(a_np, -1))
        assert orc.cholesky_residual(a_np, np.triu(out)) <= TOL_RESIDUAL_PER_N * n
    finally:
        eng.set_potrf_schedule(0); eng.set_lookahead(16, -1)


def test_potrf_lookahead_info(blocks):
    eng = blocks
    eng.set_block_sizes(256, 512); eng.set_potrf_schedule(0)
    try:
        n = 900
        rng = np.random.default_rng(4)
        p = rng.random((2 * n, n)) - 0.5
        s = p.T @ p + 0.5 * np.eye(n)
        A = dev(s); eng.reset_info(); eng._potrf(A, n, "U")
        assert eng.info() == 0 and relerr(np.triu(host(A)), np.linalg.cholesky(s).T) <= 1e-10
        s[600, 600] = -1.0
        A = dev(s); eng.reset_info(); eng._potrf(A, n, "U")
        assert eng.info() == 601
        eng.reset_info()
    finally:
        eng.set_potrf_schedule(0)


@pytest.mark.parametrize("world,n,nb,pipeline", [(2, 1500, 256, 1), (2, 4096, 512, 0), (4, 3000, 256, 1), (8, 4096, 256, 1),
                                                   (2, 1500, 256, -1), (2, 4096, 128, -1), (4, 3000, 256, -1), (8, 4096, 256, -1), (8, 4000, 128, -1)])
def test_potrf_1d_multi_gpu(world, n, nb, pipeline):
    import os
    import subprocess
    import sys
    from conftest import ROOT
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world, "--master-addr", "127.0.0.1",
           "--master-port", "29561", os.path.join(ROOT, "tools", "potrf1d_check.py"), str(n), str(nb), "--pipeline", str(pipeline)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.count("OK") == world, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.parametrize("n", [1500, 3000, 5000])
def test_potrf_default_schedule_vs_numpy(engine, n):
    """cap_dpotrf as shipped (block size by n, look-ahead schedule) -- the lapack::engine::_potrf a caller gets."""
    from capital_b200.engine import colmajor
    A = colmajor(n, n); engine.fill_symmetric(A, 0, 0, 1, 1, n, True)
    a_np = orc.symmetric_global(n)
    engine.reset_info(); engine._potrf(A, n, "U")
    out = host(A)
    assert engine.info() == 0
    assert relerr(np.triu(out), np.linalg.cholesky(a_np).T) <= TOL_FACTOR
    assert np.array_equal(np.tril(out, -1), np.tril(a_np, -1))
