"""Multi-GPU parity of the process-grid algorithms (SUMMA gemm/trmm/syrk, distributed cholinv) against numpy on
the assembled global matrices.  torchrun --nproc-per-node P tools/grid_check.py [N]   (P in 1, 2, 4, 8)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from capital_b200 import dist as cdist
from capital_b200.engine import Engine, colmajor, to_colmajor


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    eng = Engine(lr)
    comm = cdist.Comm(eng) if world > 1 else None
    c = cdist.default_depth(world)
    grid = cdist.Grid(eng, comm, c)
    d, x, y, z = grid.d, grid.x, grid.y, grid.z
    assert (x, y, z) == cdist.grid_coords(rank, c, d)
    rng = np.random.default_rng(7)
    A = rng.random((N, N)); B = rng.random((N, N)); C0 = rng.random((N, N))
    T = np.triu(rng.random((N, N))) + N * np.eye(N)
    S = rng.random((N, N)); S = S + S.T + 2 * N * np.eye(N)
    loc = lambda g: to_colmajor(np.ascontiguousarray(cdist.cyclic_block(g, x, y, d)))
    l = N // d
    ok = True

    def gather(t):
        """assemble the global matrix from layer z == 0 on every rank"""
        flat = t.t().contiguous()          # (cols, rows) row-major == the column-major storage of t
        parts = [torch.empty_like(flat) for _ in range(world)] if world > 1 else [flat]
        if world > 1:
            dist.all_gather(parts, flat)
        blocks = {}
        for r in range(world):
            xx, yy, zz = cdist.grid_coords(r, c, d)
            if zz == 0:
                blocks[(xx, yy)] = parts[r].t().cpu().numpy()
        return cdist.assemble_cyclic(blocks, d)

    def report(name, got, ref, tol=1e-11):
        nonlocal ok
        err = float(np.linalg.norm(got - ref) / np.linalg.norm(ref))
        good = err <= tol
        ok = ok and good
        if rank == 0:
            print("%-34s rel err %.2e %s" % (name, err, "ok" if good else "FAIL"), flush=True)

    # gemm
    a, b, cm = loc(A), loc(B), loc(C0)
    eng.summa_gemm(grid, a, b, cm, l, l, l, "N", "N", 1.5, 0.5); torch.cuda.synchronize()
    report("summa gemm NN alpha beta", gather(cm), 1.5 * A @ B + 0.5 * C0)
    # trmm left / right (upper, no-trans): local blocks of an upper-triangular global matrix
    t, b = loc(T), loc(B)
    eng.summa_trmm(grid, t, b, l, l, "L", "U", "N", "N", 1.0); torch.cuda.synchronize()
    report("summa trmm L U N", gather(b), T @ B)
    b = loc(B)
    eng.summa_trmm(grid, t, b, l, l, "R", "U", "N", "N", -1.0); torch.cuda.synchronize()
    report("summa trmm R U N", gather(b), -B @ T)
    # trmm left upper trans needs the grid-transposed triangular block (as cholinv.hpp:116-118)
    tt = colmajor(l, l); eng.grid_transpose(grid, t, tt)
    b = loc(B)
    # note: t is column-major; its storage is sent as is, the partner's block arrives in the same layout
    eng.summa_trmm(grid, tt, b, l, l, "L", "U", "T", "N", 1.0); torch.cuda.synchronize()
    report("summa trmm L U T (transposed T)", gather(b), T.T @ B)
    # syrk: C(upper) = -A^T A + C
    a, cm = loc(A), loc(S)
    eng.summa_syrk(grid, a, cm, l, l, "U", "T", -1.0, 1.0); torch.cuda.synchronize()
    report("summa syrk U T", np.triu(gather(cm)), np.triu(S - A.T @ A))
    # cholinv
    for ci in (True, False):
        s_loc, xinv = loc(S), colmajor(l, l)
        eng.reset_info()
        eng.cholinv_factor(grid, s_loc, xinv, ci, 1, max(256, N // 4)); torch.cuda.synchronize()
        R = np.linalg.cholesky(S).T
        report("cholinv R  (complete_inv=%d)" % ci, gather(s_loc), R)
        Xg, Xr = gather(xinv), np.linalg.inv(R)
        if not ci:   # the top-level off-diagonal block of the inverse is left unset (cholinv.hpp:147)
            h2 = (l >> 1) * d
            perm = np.concatenate([np.arange(N).reshape(-1, d)[: l >> 1].ravel(), np.arange(N).reshape(-1, d)[l >> 1:].ravel()])
            Xg = Xg.copy(); Xr = Xr.copy()
            top = np.arange(N).reshape(-1, d)[: l >> 1].ravel(); bot = np.arange(N).reshape(-1, d)[l >> 1:].ravel()
            Xg[np.ix_(top, bot)] = 0; Xr[np.ix_(top, bot)] = 0
        report("cholinv Rinv (complete_inv=%d)" % ci, Xg, Xr)
        assert eng.info() == 0
        if ci:
            # device validator (cholesky::validate::residual) against the same metric in numpy, on a clean and on a
            # deliberately damaged factor (so that agreement is not just "both tiny")
            for damage in (False, True):
                if damage and rank == 0:
                    s_loc[0, 0] *= 1.5
                got = eng.validate_cholinv(grid, loc(S), s_loc)
                Rg = gather(s_loc)
                ref = float(np.linalg.norm(np.triu(Rg.T @ Rg - S)) / np.linalg.norm(np.triu(S)))
                good = (abs(got - ref) <= 1e-8 * ref) if damage else (got <= 1e-14 and ref <= 1e-14)
                ok = ok and good
                if rank == 0:
                    print("%-34s device %.3e numpy %.3e %s" % ("validate cholinv%s" % (" (damaged)" if damage else ""), got, ref, "ok" if good else "FAIL"), flush=True)
    # 3D CholeskyQR2 on the grid against the QR of the assembled global matrix (R unique: positive diagonal)
    M = 8 * N
    Atall = rng.random((M, N))
    a_loc = loc(Atall); ml = M // d
    for iters in (1, 2):
        q_loc, r_loc = colmajor(ml, l), colmajor(l, l)
        eng.reset_info()
        eng.cacqr_factor_3d(grid, a_loc, q_loc, r_loc, iters, None, 1, max(256, N // 4)); torch.cuda.synchronize()
        Qr, Rr = np.linalg.qr(Atall)
        sg = np.sign(np.diag(Rr)); Qr = Qr * sg; Rr = (Rr.T * sg).T
        Qg, Rg = gather(q_loc), gather(r_loc)
        report("3D CQR%d R" % iters, Rg, Rr, 1e-11 if iters == 2 else 1e-9)
        report("3D CQR%d Q" % iters, Qg, Qr, 1e-11 if iters == 2 else 1e-9)
        orth = float(np.linalg.norm(Qg.T @ Qg - np.eye(N)))
        good = orth <= (1e-13 * N if iters == 2 else 1e-9)
        ok = ok and good
        if rank == 0:
            print("%-34s %.2e %s" % ("3D CQR%d ||Q^T Q - I||_F" % iters, orth, "ok" if good else "FAIL"), flush=True)
        assert eng.info() == 0
        if iters == 2:
            for damage in (False, True):
                if damage and rank == 0:
                    q_loc[0, 0] += 0.25
                o_dev, r_dev = eng.validate_cacqr_3d(grid, a_loc, q_loc, r_loc)
                Qg = gather(q_loc)
                o_ref = float(np.linalg.norm(Qg.T @ Qg - np.eye(N)) / N)
                r_ref = float(np.linalg.norm(Qg @ Rg - Atall) / np.linalg.norm(Atall))
                good = (abs(o_dev - o_ref) <= 1e-8 * o_ref and abs(r_dev - r_ref) <= 1e-8 * r_ref) if damage else (o_dev <= 1e-14 and r_dev <= 1e-14)
                ok = ok and good
                if rank == 0:
                    print("%-34s orth %.3e / %.3e  resid %.3e / %.3e %s" % ("validate 3D CQR2%s" % (" (damaged)" if damage else ""), o_dev, o_ref, r_dev, r_ref, "ok" if good else "FAIL"), flush=True)
    if rank == 0:
        print("GRID CHECK", "PASSED" if ok else "FAILED", "world", world, "grid %dx%dx%d" % (c, d, d), "N", N, flush=True)
    grid.close()
    if comm:
        comm.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
