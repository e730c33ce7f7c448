"""GPU parity: the CUDA path, called through the C ABI (ctypes -> libcapital_b200.so), against
(1) fixtures from the unmodified reference (tests/golden), (2) the numpy oracle on seeded inputs,
(3) size-independent properties at BASELINE sizes.  Tolerances: BASELINE.json north_star
(relative factor difference <= 1e-11; residual / orthogonality <= 1e-13 * n)."""
import numpy as np
import pytest
import torch

import capital_oracle as orc
from conftest import TOL_FACTOR, TOL_RESIDUAL_PER_N, golden, relerr

pytestmark = pytest.mark.gpu


def dev(a):
    from capital_b200.engine import to_colmajor
    return to_colmajor(np.asarray(a))


def host(t):
    torch.cuda.synchronize()
    return t.cpu().numpy()


@pytest.mark.parametrize("ta,tb", [("N", "N"), ("T", "N"), ("N", "T"), ("T", "T")])
def test_gemm_vs_reference_engine(engine, ta, tb):
    e = golden("engines"); k = "gemm_%s%s_" % (ta, tb)
    A, B, Cm = dev(e[k + "A"]), dev(e[k + "B"]), dev(e[k + "C0"])
    m, n = e[k + "C"].shape; kk = e[k + "A"].shape[0] if ta == "T" else e[k + "A"].shape[1]
    engine._gemm(A, B, Cm, m, n, kk, ta, tb, 1.25, -0.5)
    assert relerr(host(Cm), e[k + "C"]) <= 1e-13


def test_syrk_vs_reference_engine(engine):
    e = golden("engines")
    A, Cm = dev(e["syrk_A"]), dev(e["syrk_C0"]); k, n = e["syrk_A"].shape
    engine._syrk(A, Cm, n, k, "U", "T", 1.0, 0.0)
    out = host(Cm)
    assert relerr(np.triu(out), np.triu(e["syrk_C"])) <= 1e-13
    assert np.array_equal(np.tril(out, -1), np.tril(e["syrk_C0"], -1))       # strict lower untouched, like dsyrk


@pytest.mark.parametrize("side,uplo,tr", [("R", "U", "N"), ("L", "U", "T"), ("L", "U", "N"), ("R", "L", "T")])
def test_trmm_vs_reference_engine(engine, side, uplo, tr):
    e = golden("engines"); k = "trmm_%s%s%s_" % (side, uplo, tr)
    T, B = dev(e[k + "T"]), dev(e[k + "B0"]); m, n = e[k + "B0"].shape
    engine._trmm(T, B, m, n, side, uplo, tr, "N", -1.0)
    assert relerr(host(B), e[k + "B"]) <= 1e-13


def test_potrf_trtri_vs_reference_engine(engine):
    e = golden("engines"); n = e["potrf_A"].shape[0]
    A = dev(e["potrf_A"]); engine.reset_info(); engine._potrf(A, n, "U")
    out = host(A)
    assert relerr(np.triu(out), np.triu(e["potrf_R"])) <= TOL_FACTOR
    assert np.array_equal(np.tril(out, -1), np.tril(e["potrf_A"], -1))       # strict lower untouched, like dpotrf
    assert engine.info() == 0
    T = dev(e["trtri_T"]); engine._trtri(T, n, "U", "N")
    out = host(T)
    assert relerr(np.triu(out), np.triu(e["trtri_X"])) <= TOL_FACTOR
    assert np.array_equal(np.tril(out, -1), np.tril(e["trtri_T"], -1))


def test_potrf_reports_non_spd(engine):
    A = dev(np.array([[4.0, 2.0, 1.0], [2.0, 1.0, 3.0], [1.0, 3.0, 1.0]]))   # second leading minor is singular
    engine.reset_info(); engine._potrf(A, 3, "U")
    assert engine.info() == 2
    engine.reset_info()


def test_generators_bit_exact(engine):
    from capital_b200.engine import colmajor
    g = golden("generators")
    for ranks, x, y in ((1, 7, 20), (4, 6, 32)):
        for r in range(ranks):
            ref = g["random_p%d_x%d_y%d_r%d" % (ranks, x, y, r)]
            A = colmajor(*ref.shape); engine.fill_random(A, r)
            assert np.array_equal(host(A), ref)
    for ranks, n in ((1, 9), (4, 10)):
        d = int(round(ranks ** 0.5))
        for r in range(ranks):
            ref = g["symmetric_p%d_n%d_r%d" % (ranks, n, r)]
            A = colmajor(*ref.shape); engine.fill_symmetric(A, r // d, r % d, d, d, n, True)
            assert np.array_equal(host(A), ref)
    big = colmajor(100003, 3); engine.fill_random(big, 77)                   # jump-ahead across thread chunks
    assert np.array_equal(host(big), orc.distribute_random(3, 100003, 0, 0, 1, 1, 77))


@pytest.mark.parametrize("tag", ["cqr2_p1", "cqr2_p4", "cqr1_p2"])
def test_cacqr_vs_reference(engine, tag):
    """Multi-rank fixtures are replayed on one GPU: per-rank Gram, rank-ordered sum (the allreduce),
    redundant potrf+trtri, per-rank Q = A*Rinv -- the schedule of cacqr.hpp:7-29."""
    from capital_b200.engine import colmajor
    q = golden("cacqr"); ranks, variant, m, n = [int(v) for v in q[tag + "_meta"]]
    blocks = [dev(q["%s_A_r%d" % (tag, r)]) for r in range(ranks)]
    qs = [colmajor(m // ranks, n) for _ in range(ranks)]
    r_total = None
    src = blocks
    for sweep in range(variant):
        G = colmajor(n, n, fill=0.0); part = colmajor(n, n, fill=0.0)
        for r in range(ranks):
            engine.cqr_gram(src[r], part); G += torch.triu(part)
        X = colmajor(n, n); engine.cholinv(G, X, n, True)
        for r in range(ranks):
            engine.cqr_apply(X, src[r], qs[r])
        Rk = torch.triu(G).clone()
        r_total = Rk if r_total is None else Rk @ r_total
        src = qs
    for r in range(ranks):
        assert relerr(host(qs[r]), q["%s_Q_r%d" % (tag, r)]) <= TOL_FACTOR
    if variant == 2:
        assert relerr(host(r_total), q[tag + "_R"]) <= TOL_FACTOR
    if ranks == 1:      # the fused entry point
        Q = colmajor(m, n); R = colmajor(n, n); engine.cacqr_factor_1d(blocks[0], Q, R, variant)
        assert relerr(host(Q), q[tag + "_Q_r0"]) <= TOL_FACTOR
        assert relerr(host(R), q[tag + "_R"]) <= TOL_FACTOR
        assert np.count_nonzero(np.tril(host(R), -1)) == 0


@pytest.mark.parametrize("tag", ["p1_ci1", "p1_ci0", "p8_ci1"])
def test_cholinv_vs_reference(engine, tag):
    from capital_b200.engine import colmajor
    c = golden("cholinv"); ranks, n, ci, bc, d = [int(v) for v in c[tag + "_meta"]]
    cdim = 1 if ranks == 1 else 2
    def assemble(name):
        blocks = {}
        for r in range(ranks):
            z, y, x = r % cdim, r // (d * cdim), (r % (d * cdim)) // cdim
            if z == 0:
                blocks[(x, y)] = c["%s_%s_r%d" % (tag, name, r)]
        return orc.assemble_cyclic_2d(blocks, n, n, d)
    A = colmajor(n, n); engine.fill_symmetric(A, 0, 0, 1, 1, n, True)
    assert np.array_equal(host(A), assemble("A"))
    X = colmajor(n, n); engine.reset_info(); engine.cholinv(A, X, n, bool(ci))
    assert relerr(np.triu(host(A)), np.triu(assemble("R"))) <= TOL_FACTOR
    if ci:
        assert relerr(host(X), np.triu(assemble("Rinv"))) <= TOL_FACTOR
    assert engine.info() == 0


@pytest.mark.parametrize("tag", ["p1", "p8"])
def test_summa_gemm_vs_reference(engine, tag):
    from capital_b200.engine import colmajor
    s = golden("summa"); ranks, M, N, K, d = [int(v) for v in s[tag + "_meta"]]
    cdim = 1 if ranks == 1 else 2
    def assemble(name, rows, cols):
        blocks = {}
        for r in range(ranks):
            z, y, x = r % cdim, r // (d * cdim), (r % (d * cdim)) // cdim
            if z == 0:
                blocks[(x, y)] = s["%s_%s_r%d" % (tag, name, r)]
        return orc.assemble_cyclic_2d(blocks, rows, cols, d)
    A, B = dev(assemble("A", M, K)), dev(assemble("B", K, N)); Cm = colmajor(M, N)
    engine._gemm(A, B, Cm, M, N, K)
    assert relerr(host(Cm), assemble("C", M, N)) <= 1e-13


@pytest.mark.parametrize("m,n", [(1, 1), (5, 3), (130, 129), (1000, 257), (4096, 256), (65536, 256), (20000, 1024)])
def test_cacqr2_vs_oracle_seeded(engine, m, n):
    """Seeded drand48 inputs (the bench generator) at sizes the oracle finishes in seconds; includes
    ragged shapes (m, n not multiples of the 128 tile) and cfg1's 65536 x 256."""
    from capital_b200.engine import colmajor
    A = colmajor(m, n); engine.fill_random(A, 3); Q = colmajor(m, n); R = colmajor(n, n)
    engine.reset_info(); engine.cacqr_factor_1d(A, Q, R, 2)
    a_np = orc.distribute_random(n, m, 0, 0, 1, 1, 3)
    if m >= n:
        qo, ro = orc.cacqr_factor_1d([a_np], 2)
        assert relerr(host(R), ro) <= TOL_FACTOR and relerr(host(Q), qo[0]) <= TOL_FACTOR
        assert orc.qr_orthogonality(host(Q)) <= TOL_RESIDUAL_PER_N * n
        assert orc.qr_residual(a_np, host(Q), host(R)) <= TOL_RESIDUAL_PER_N * n
        assert engine.info() == 0


def test_cacqr_empty_block(engine):
    from capital_b200.engine import colmajor
    n = 8
    G = colmajor(n, n, fill=5.0)
    engine.cqr_gram(colmajor(0, n), G)          # a rank that owns no rows contributes a zero Gram
    assert np.count_nonzero(np.triu(host(G))) == 0


@pytest.mark.parametrize("n", [1, 2, 127, 128, 129, 300, 1024, 1500])
def test_cholinv_vs_oracle_seeded(engine, n):
    from capital_b200.engine import colmajor
    A = colmajor(n, n); engine.fill_symmetric(A, 0, 0, 1, 1, n, True); a_np = orc.symmetric_global(n)
    X = colmajor(n, n); engine.cholinv(A, X, n, True)
    R, Rinv = orc.cholinv_factor(a_np, True, 1, 128)
    assert relerr(np.triu(host(A)), R) <= TOL_FACTOR and relerr(host(X), Rinv) <= TOL_FACTOR
    assert orc.cholesky_residual(a_np, np.triu(host(A))) <= TOL_RESIDUAL_PER_N * n


def test_trtri_lower_and_roundtrip(engine):
    n = 777
    rng = np.random.default_rng(5)
    L = np.tril(rng.random((n, n))) + n * np.eye(n) + np.triu(rng.random((n, n)), 1)     # garbage above the diagonal
    T = dev(L); engine._trtri(T, n, "L", "N")
    out = host(T)
    assert relerr(np.tril(out), np.linalg.inv(np.tril(L))) <= TOL_FACTOR
    assert np.array_equal(np.triu(out, 1), np.triu(L, 1))
    engine._trtri(T, n, "L", "N")                                                         # inverse of the inverse
    assert relerr(np.tril(host(T)), np.tril(L)) <= 1e-10


def test_pack_unpack_axpby(engine):
    from capital_b200.engine import colmajor
    n = 131
    rng = np.random.default_rng(1); F = rng.random((n, n))
    full = dev(F); packed = torch.empty(n * (n + 1) // 2, dtype=torch.float64, device="cuda")
    engine.pack_upper(packed, full, n)
    ref = np.concatenate([F[: c + 1, c] for c in range(n)])                 # uppertri offset X(X+1)/2+Y (structure.h:39)
    assert np.array_equal(host(packed), ref)
    back = colmajor(n, n, fill=9.0); engine.unpack_upper(back, packed, n, True)
    assert np.array_equal(host(back), np.triu(F))
    y = dev(F); x = dev(2 * F); engine.axpby(y, x, 0.5, 2.0)               # C = beta*C + S of summa.hpp:33
    assert relerr(host(y), 3 * F) <= 1e-15


def test_cacqr2_full_baseline_size_properties(engine):
    """BASELINE cfg2 shape on one GPU (2^22 x 1024, 32 GiB): properties only -- orthogonality via an
    independent torch/cuBLAS product, R upper with positive diagonal, and A = QR on a row sample."""
    from capital_b200.engine import colmajor
    m, n = 1 << 22, 1024
    free, _ = torch.cuda.mem_get_info()
    if free < 80 * (1 << 30):
        pytest.skip("needs 80 GB of free HBM")
    A = colmajor(m, n); engine.fill_random(A, 0); Q = colmajor(m, n); R = colmajor(n, n)
    engine.reset_info(); engine.cacqr_factor_1d(A, Q, R, 2); torch.cuda.synchronize()
    assert engine.info() == 0
    G = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    step = 1 << 19
    for s in range(0, m, step):
        blk = Q[s:s + step]
        G += blk.t() @ blk
    orth = (torch.linalg.norm(G - torch.eye(n, dtype=torch.float64, device="cuda")) / n).item()
    assert orth <= TOL_RESIDUAL_PER_N * n
    Rh = host(R)
    assert np.count_nonzero(np.tril(Rh, -1)) == 0 and np.all(np.diag(Rh) > 0)
    rows = torch.arange(0, m, 4099, device="cuda")
    res = (torch.linalg.norm(Q[rows] @ R - A[rows]) / torch.linalg.norm(A[rows])).item()
    assert res <= TOL_RESIDUAL_PER_N * n


@pytest.mark.parametrize("num_iter", [1, 2])
def test_cacqr_host_entry_matches_device_path(engine, num_iter):
    """cap_cacqr_factor_1d_host (pinned host buffers, chunked two-stream pipeline) against the device-resident call;
    the chunk size is shrunk so that the 5000-row block is cut into several ragged chunks."""
    import torch
    from capital_b200.engine import colmajor
    m, n = 5000, 96
    A = colmajor(m, n); engine.fill_random(A, 5)
    Q = colmajor(m, n); R = colmajor(n, n)
    engine.cacqr_factor_1d(A, Q, R, num_iter); torch.cuda.synchronize()
    hA = torch.empty((n, m), dtype=torch.float64, pin_memory=True); hA.copy_(A.t())
    hQ = torch.empty((n, m), dtype=torch.float64, pin_memory=True); hR = torch.empty((n, n), dtype=torch.float64, pin_memory=True)
    engine.set_host_chunk_bytes(1 << 20)
    try:
        engine.cacqr_factor_1d_host(hA, hQ, hR, m, n, num_iter)
    finally:
        engine.set_host_chunk_bytes(1 << 30)
    assert relerr(hQ.t().numpy(), host(Q)) <= 1e-13
    assert relerr(np.triu(hR.t().numpy()), np.triu(host(R))) <= 1e-13
    assert orc.qr_orthogonality(hQ.t().numpy()) <= TOL_RESIDUAL_PER_N * n * (1 if num_iter == 2 else 1e3)


@pytest.mark.gpu
@pytest.mark.parametrize("m,n", [(1000, 257), (65536, 256)])
def test_device_validators_match_oracle_metrics(engine, m, n):
    """qr::validate::{orthogonality,residual} on the device (cap_validate_cacqr_1d) against the oracle's restatement
    of the same formulas: tiny on a clean factorization, and equal to 1e-8 relative on a deliberately damaged Q."""
    from capital_b200.engine import colmajor
    A = colmajor(m, n); engine.fill_random(A, 5); Q = colmajor(m, n); R = colmajor(n, n)
    engine.cacqr_factor_1d(A, Q, R, 2)
    orth, resid = engine.validate_cacqr_1d(A, Q, R)
    assert orth <= TOL_RESIDUAL_PER_N * n and resid <= TOL_RESIDUAL_PER_N * n
    assert abs(orth - orc.qr_orthogonality(host(Q))) <= 1e-15 and abs(resid - orc.qr_residual(host(A), host(Q), host(R))) <= 1e-15
    Q[3, 2] += 0.5
    orth, resid = engine.validate_cacqr_1d(A, Q, R)
    o_ref, r_ref = orc.qr_orthogonality(host(Q)), orc.qr_residual(host(A), host(Q), host(R))
    assert abs(orth - o_ref) <= 1e-8 * o_ref and abs(resid - r_ref) <= 1e-8 * r_ref
