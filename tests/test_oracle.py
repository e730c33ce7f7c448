"""CPU: the numpy restatement (oracle/capital_oracle.py) against fixtures produced by the unmodified
reference compiled here (tests/golden/make_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest

import capital_oracle as orc
from conftest import TOL_FACTOR, golden, relerr


def test_generators_bit_exact():
    g = golden("generators")
    for ranks, x, y in ((1, 7, 20), (3, 5, 10), (4, 6, 32)):
        for r in range(ranks):
            mine = orc.distribute_random(x, y, 0, r, 1, ranks, r)
            assert np.array_equal(mine, g["random_p%d_x%d_y%d_r%d" % (ranks, x, y, r)])
    for ranks, n in ((1, 9), (4, 10), (4, 7)):
        d = int(round(ranks ** 0.5))
        for r in range(ranks):
            mine = orc.distribute_symmetric(n, n, r // d, r % d, d, d, True)
            assert np.array_equal(mine, g["symmetric_p%d_n%d_r%d" % (ranks, n, r)])


def test_drand48_jump_ahead_matches_stream():
    s = orc.drand48_stream(12345, 70000)
    a_t, c_t = orc._affine_pow(65537)
    x = (a_t * orc.lcg_seed(12345) + c_t) & ((1 << 48) - 1)
    assert s[65536] == x / float(1 << 48)
    assert np.all((s >= 0) & (s < 1))


@pytest.mark.parametrize("tag", ["cqr2_p1", "cqr2_p4", "cqr1_p2"])
def test_cacqr_matches_reference(tag):
    q = golden("cacqr")
    ranks, variant, m, n = [int(v) for v in q[tag + "_meta"]]
    blocks = orc.cacqr_input_blocks(m, n, ranks)
    for r in range(ranks):
        assert np.array_equal(blocks[r], q["%s_A_r%d" % (tag, r)])   # same inputs, bit for bit
    qb, R = orc.cacqr_factor_1d(blocks, variant)
    if variant == 2:
        assert relerr(R, q[tag + "_R"]) <= TOL_FACTOR
    else:
        # reference quirk: with num_iter=1 the Serialize policy never writes R back (complete_1d is only
        # reached for num_iter>1, cacqr.hpp:181-189), so construct_R returns the packed GRAM matrix.
        assert relerr(np.triu(R.T @ R), q[tag + "_R"]) <= 1e-13
    for r in range(ranks):
        assert relerr(qb[r], q["%s_Q_r%d" % (tag, r)]) <= TOL_FACTOR
    A = orc.assemble_row_cyclic(blocks, m); Q = orc.assemble_row_cyclic(qb, m)
    assert orc.qr_residual(A, Q, R) <= 1e-13 * n and orc.qr_orthogonality(Q) <= 1e-13 * n


@pytest.mark.parametrize("tag", ["ci1", "ci0"])
def test_cacqr_3d_reference_fixture(tag):
    """tests/golden/cacqr3d.npz = the reference's 3D CholeskyQR2 on the 2x2x2 cube (topo::rect c = 2, 8 mini-MPI ranks;
    ci0 goes through cacqr::solve).  Q and R of a full-rank matrix are unique, so the 1D restatement on the assembled matrix
    must reproduce them; the inputs are the generator's (key = rank / c), bit for bit, and replicated over depth."""
    from capital_b200 import dist as cdist
    q = golden("cacqr3d")
    ranks, variant, m, n, c, d, ci = [int(v) for v in q[tag + "_meta"]]
    A, Q, R = {}, {}, {}
    for r in range(ranks):
        x, y, z = cdist.grid_coords(r, c, d)                       # topo::rect (topology.h:47-49) = z fastest, then x, then y
        assert np.array_equal(q["%s_A_r%d" % (tag, r)], orc.distribute_random(n, m, x, y, c, d, r // c))
        if z == 0:
            A[(x, y)], Q[(x, y)], R[(x, y)] = (q["%s_%s_r%d" % (tag, k, r)] for k in "AQR")
        else:
            for k in "AQR":
                assert np.array_equal(q["%s_%s_r%d" % (tag, k, r)], q["%s_%s_r%d" % (tag, k, r - z)])
    Ag, Qg, Rg = (cdist.assemble_cyclic(b, d) for b in (A, Q, R))
    rows = [Ag[p::4] for p in range(4)]                          # any row deal gives the same factors
    qb, R1 = orc.cacqr_factor_1d(rows, 2)
    Q1 = np.empty_like(Ag)
    for p in range(4):
        Q1[p::4] = qb[p]
    assert relerr(np.triu(Rg), R1) <= TOL_FACTOR and relerr(Qg, Q1) <= TOL_FACTOR
    assert np.all(np.tril(Rg, -1) == 0)
    assert orc.qr_residual(Ag, Qg, np.triu(Rg)) <= 1e-13 * n and orc.qr_orthogonality(Qg) <= 1e-13 * n


@pytest.mark.parametrize("tag", ["p1_ci1", "p1_ci0", "p8_ci1", "p8_ci0"])
def test_cholinv_matches_reference(tag):
    c = golden("cholinv")
    ranks, n, ci, bc, d = [int(v) for v in c[tag + "_meta"]]
    cdim = 1 if ranks == 1 else 2
    def assemble(name):
        blocks = {}
        for r in range(ranks):
            z, y, x = r % cdim, r // (d * cdim), (r % (d * cdim)) // cdim      # topo::square layout 0 (topology.h:80-83)
            if z == 0:
                blocks[(x, y)] = c["%s_%s_r%d" % (tag, name, r)]
        return orc.assemble_cyclic_2d(blocks, n, n, d)
    A = assemble("A")
    assert np.array_equal(A, orc.symmetric_global(n))
    R, Rinv = orc.cholinv_factor(A, bool(ci), 1, orc.cholinv_bc_dimension(n, cdim, d, bc))
    assert relerr(R, np.triu(assemble("R"))) <= TOL_FACTOR
    Rinv_ref = np.triu(assemble("Rinv"))
    if ci:
        assert relerr(Rinv, Rinv_ref) <= TOL_FACTOR
    else:   # top-level off-diagonal block of Rinv is never formed (cholinv.hpp:147)
        h = n // 2
        assert relerr(Rinv[:h, :h], Rinv_ref[:h, :h]) <= TOL_FACTOR and relerr(Rinv[h:, h:], Rinv_ref[h:, h:]) <= TOL_FACTOR
    assert orc.cholesky_residual(A, R) <= 1e-13 * n


@pytest.mark.parametrize("tag", ["p1", "p8"])
def test_summa_matches_reference(tag):
    s = golden("summa")
    ranks, M, N, K, d = [int(v) for v in s[tag + "_meta"]]
    cdim = 1 if ranks == 1 else 2
    def assemble(name, rows, cols):
        blocks = {}
        for r in range(ranks):
            z, y, x = r % cdim, r // (d * cdim), (r % (d * cdim)) // cdim
            if z == 0:
                blocks[(x, y)] = s["%s_%s_r%d" % (tag, name, r)]
        return orc.assemble_cyclic_2d(blocks, rows, cols, d)
    A, B, Cm = assemble("A", M, K), assemble("B", K, N), assemble("C", M, N)
    assert relerr(orc.summa_gemm(A, B), Cm) <= 1e-13


def test_engine_restatements_match_openblas():
    e = golden("engines")
    assert relerr(orc.syrk_upper_trans(e["syrk_A"]), np.triu(e["syrk_C"])) <= 1e-13
    assert np.array_equal(np.tril(e["syrk_C"], -1), np.tril(e["syrk_C0"], -1))     # strict lower untouched
    assert relerr(orc.potrf_upper(e["potrf_A"]), np.triu(e["potrf_R"])) <= TOL_FACTOR
    assert relerr(orc.trtri_upper(e["trtri_T"]), np.triu(e["trtri_X"])) <= TOL_FACTOR
    assert relerr(orc.trmm_right_upper(e["trmm_RUN_B0"], e["trmm_RUN_T"]) * -1.0, e["trmm_RUN_B"]) <= 1e-13


def test_potrf_rejects_non_spd():
    with pytest.raises(np.linalg.LinAlgError):
        orc.potrf_upper(np.array([[1.0, 2.0], [2.0, 1.0]]))
