"""GPU: process-grid algorithms (SUMMA gemm/trmm/syrk, distributed cholinv) through the C ABI against numpy on the
assembled global matrices (tools/grid_check.py), at every world size the box offers; the C++ host drivers; and
the two tile engines against each other."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, TOL_FACTOR, relerr

pytestmark = pytest.mark.gpu


def _run_grid_check(world, n, port):
    script = os.path.join(ROOT, "tools", "grid_check.py")
    if world == 1:
        cmd = [sys.executable, script, str(n)]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world, "--master-addr", "127.0.0.1",
               "--master-port", str(port), script, str(n)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "GRID CHECK PASSED" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.parametrize("world,n", [(1, 640), (2, 1024), (4, 1024), (8, 2048)])
def test_grid_algorithms(world, n):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    _run_grid_check(world, n, 29540 + world)


@pytest.mark.parametrize("tag", ["ci1", "ci0"])
def test_cacqr_3d_vs_reference_fixture(engine, tag):
    """cap_cacqr_factor_3d against the REFERENCE's 3D CholeskyQR2 (tests/golden/cacqr3d.npz: qr::cacqr::factor on the 2x2x2
    cube, unmodified reference on 8 mini-MPI ranks; ci0 = cacqr::solve).  On one GPU the grid is 1x1x1 and the local block is
    the assembled global matrix; Q and R are unique, so they must agree with the reference's assembled factors.  (At 8 ranks
    capital_b200.checks.grid_small_check compares every rank's local blocks with the fixture directly.)"""
    import torch
    from capital_b200 import dist as cdist
    from capital_b200.engine import colmajor, to_colmajor
    from conftest import golden
    q = golden("cacqr3d")
    ranks, variant, m, n, c, d, ci = [int(v) for v in q[tag + "_meta"]]
    blocks = {k: {} for k in "AQR"}
    for r in range(ranks):
        x, y, z = cdist.grid_coords(r, c, d)
        if z == 0:
            for k in "AQR":
                blocks[k][(x, y)] = q["%s_%s_r%d" % (tag, k, r)]
    Ag, Qg, Rg = (cdist.assemble_cyclic(blocks[k], d) for k in "AQR")
    grid = cdist.Grid(engine, None, 1)
    engine.set_cacqr_complete_inv(bool(ci))
    try:
        A = to_colmajor(Ag); Q = colmajor(m, n); R = colmajor(n, n)
        engine.reset_info()
        engine.cacqr_factor_3d(grid, A, Q, R, variant, None, 1, 16); torch.cuda.synchronize()
        assert engine.info() == 0
        assert relerr(np.triu(R.cpu().numpy()), np.triu(Rg)) <= TOL_FACTOR
        assert relerr(Q.cpu().numpy(), Qg) <= TOL_FACTOR
    finally:
        engine.set_cacqr_complete_inv(True)
        grid.close()


def test_reference_engine_headers_bound_to_device(tmp_path):
    """integration/_build/engine_check = the reference's UNMODIFIED src/blas/{engine.h,interface.h} and
    src/lapack/{engine.h,interface.h} compiled (in the build container, integration/Makefile) with
    integration/{blas,lapack}/interface.hpp in place of the CBLAS / LAPACKE specialisations.  Operands and expected
    results: tests/golden/engines.npz, the outputs of the reference's own engines on the CPU."""
    exe = os.path.join(ROOT, "integration", "_build", "engine_check")
    if not os.path.exists(exe):
        pytest.skip("integration/_build/engine_check not built (make -C integration, needs /root/reference)")
    z = np.load(os.path.join(ROOT, "tests", "golden", "engines.npz"))
    put = lambda name, a: np.asfortranarray(a).T.copy().tofile(str(tmp_path / name))      # column-major raw doubles
    lines, expect = [], {}
    for tag in ("NN", "TN", "NT", "TT"):
        A, B, C0, C = (z["gemm_%s_%s" % (tag, k)] for k in ("A", "B", "C0", "C"))
        m, n = C.shape; k = A.shape[0] if tag[0] == "T" else A.shape[1]
        put("g%sA" % tag, A); put("g%sB" % tag, B); put("g%sC" % tag, C0)
        lines.append("gemm %s %s %d %d %d 1.25 -0.5 g%sA g%sB g%sC g%sO" % (tag[0], tag[1], m, n, k, tag, tag, tag, tag))
        expect["g%sO" % tag] = (C, None)
    A, C0, C = z["syrk_A"], z["syrk_C0"], z["syrk_C"]
    put("sA", A); put("sC", C0)
    lines.append("syrk U T %d %d 1.0 0.0 sA sC sO" % (C.shape[0], A.shape[0])); expect["sO"] = (C, "U")
    for tag in ("RUN", "LUT", "LUN", "RLT"):
        T, B0, B = z["trmm_%s_T" % tag], z["trmm_%s_B0" % tag], z["trmm_%s_B" % tag]
        put("t%sT" % tag, T); put("t%sB" % tag, B0)
        lines.append("trmm %s %s %s N %d %d -1.0 t%sT t%sB t%sO" % (tag[0], tag[1], tag[2], B.shape[0], B.shape[1], tag, tag, tag)); expect["t%sO" % tag] = (B, None)
    put("pA", z["potrf_A"]); lines.append("potrf U %d pA pO" % z["potrf_A"].shape[0]); expect["pO"] = (z["potrf_R"], "U")
    put("xT", z["trtri_T"]); lines.append("trtri U N %d xT xO" % z["trtri_T"].shape[0]); expect["xO"] = (z["trtri_X"], "U")
    (tmp_path / "cases.txt").write_text("\n".join(lines) + "\n")
    out = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "%d cases" % len(lines) in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
    for name, (want, tri) in expect.items():
        got = np.fromfile(str(tmp_path / name)).reshape(want.shape[1], want.shape[0]).T
        if tri == "U":
            got, want = np.triu(got), np.triu(want)
        assert relerr(got, want) <= TOL_FACTOR, name


def test_cpp_host_drivers():
    """The reference's bench command lines (bench/qr/cacqr.cpp:14-25, bench/cholesky/cholinv.cpp:15-22,
    bench/matmult/summa_gemm.cpp:19-25) run against the C++ host mirror."""
    bindir = os.path.join(ROOT, "capital_b200", "host", "bin")
    if not os.path.exists(os.path.join(bindir, "cacqr")):
        pytest.skip("host drivers not built (python -m capital_b200.build)")
    import re
    env = dict(os.environ, CAPITAL_B200_VALIDATE="1")
    num = r"([0-9.eE+-]+)"
    for cmd, n, patterns in (
            (["cacqr", "2", "65536", "256", "1", "1", "1", "1", "0", "0", "0", "0", "1"], 256, [r"orthogonality .* - " + num + r"\s+residual", r"residual \|\|Q R - A.* - " + num]),
            (["cholinv", "2048", "1", "1", "1", "0", "0", "0", "1"], 2048, [r"residual \|\|R\^T R - A.* - " + num]),
            (["cholinv", "2048", "1", "0", "1", "-2", "1", "3", "1"], 2048, [r"residual \|\|R\^T R - A.* - " + num]),      # layout 1, num_chunks 3
            (["potrf", "3000", "256", "1"], 3000, [r"residual \|\|R\^T R - A.* - " + num]),
            (["summa_gemm", "1024", "768", "512", "1", "0", "0", "1"], 0, [])):
        out = subprocess.run([os.path.join(bindir, cmd[0])] + cmd[1:], capture_output=True, text=True, timeout=300, env=env)
        assert out.returncode == 0 and "TFLOP/s" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
        for pat in patterns:                       # the reference's validators, printed by the driver: test/qr/validate.hpp, test/cholesky/validate.hpp
            m = re.search(pat, out.stdout)
            assert m, (pat, out.stdout[-1500:])
            assert float(m.group(1)) <= 1e-13 * n, (cmd, m.group(0))


@pytest.mark.parametrize("side,uplo,tr", [(s, u, t) for s in "LR" for u in "UL" for t in "NT"])
def test_trmm_diag_skipping_both_engines(engine, side, uplo, tr):
    """Sizes that cross several 128-tiles with ragged edges: the TMA engine skips structural zeros of the
    diagonal blocks, the cp.async engine multiplies them; both must agree with numpy."""
    import torch
    from capital_b200.engine import colmajor, to_colmajor
    rng = np.random.default_rng(11)
    m, n = (520, 300) if side == "L" else (300, 520)
    nt = m if side == "L" else n
    T = rng.random((nt, nt)); B = rng.random((m, n))
    Tt = np.triu(T) if uplo == "U" else np.tril(T)
    op = Tt.T if tr == "T" else Tt
    ref = 1.25 * (op @ B if side == "L" else B @ op)
    for eng_id in (0, 1):
        engine.set_engine(eng_id)
        t, b, c = to_colmajor(T), to_colmajor(B), colmajor(m, n)
        engine.trmm_oop(t, b, c, m, n, side, uplo, tr, "N", 1.25); torch.cuda.synchronize()
        assert relerr(c.cpu().numpy(), ref) <= TOL_FACTOR, (eng_id, side, uplo, tr)
        engine._trmm(t, b, m, n, side, uplo, tr, "N", 1.25); torch.cuda.synchronize()
        assert relerr(b.cpu().numpy(), ref) <= TOL_FACTOR, ("in place", eng_id, side, uplo, tr)
    engine.set_engine(0)


@pytest.mark.parametrize("uplo,tr", [("U", "T"), ("L", "T"), ("U", "N"), ("L", "N")])
def test_syrk_masked_diagonal_tiles(engine, uplo, tr):
    import torch
    from capital_b200.engine import to_colmajor
    rng = np.random.default_rng(12)
    n, k = 650, 3000
    A = rng.random((k, n)) if tr == "T" else rng.random((n, k))
    C0 = rng.random((n, n))
    full = 0.5 * (A.T @ A if tr == "T" else A @ A.T) + 2.0 * C0
    ref = np.where(np.triu(np.ones((n, n))) > 0 if uplo == "U" else np.tril(np.ones((n, n))) > 0, full, C0)
    for eng_id in (0, 1):
        engine.set_engine(eng_id)
        a, c = to_colmajor(A), to_colmajor(C0)
        engine._syrk(a, c, n, k, uplo, tr, 0.5, 2.0); torch.cuda.synchronize()
        assert relerr(c.cpu().numpy(), ref) <= TOL_FACTOR, (eng_id, uplo, tr)
    engine.set_engine(0)


def test_unaligned_operands_take_the_cpasync_engine(engine):
    """Odd leading dimensions / 8-byte-aligned bases are outside TMA's rules: same results through the fallback."""
    import torch
    rng = np.random.default_rng(13)
    m, n, k = 257, 131, 390
    big_a = torch.as_tensor(rng.random((k + 3, m + 1)), device="cuda").t()   # column-major (m+1) x (k+3), ld = m+1 (even or odd)
    big_b = torch.as_tensor(rng.random((n + 2, k + 1)), device="cuda").t()
    A = big_a[1:m + 1, 1:k + 1]; B = big_b[1:k + 1, :n]                       # bases offset by one double
    from capital_b200.engine import colmajor
    C = colmajor(m, n)
    engine._gemm(A, B, C, m, n, k, "N", "N", 1.0, 0.0); torch.cuda.synchronize()
    assert relerr(C.cpu().numpy(), A.cpu().numpy() @ B.cpu().numpy()) <= TOL_FACTOR
