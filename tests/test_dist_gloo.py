"""CPU, world_size 2 over gloo: the host-side sharding logic of the N>1 path (row deal, per-rank
generator keys, rank-ordered Gram reduction, id broadcast, max-over-ranks timing) reproduces the
unsharded factorization.  The arithmetic here is the oracle's -- this tests plumbing, not kernels."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, m, n, out):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist
    import capital_oracle as orc
    from capital_b200 import dist as cdist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    padded, owned = cdist.local_rows(m, world, rank)
    blk = orc.distribute_random(n, m, 0, rank, 1, world, rank)            # bench/qr/cacqr.cpp:34
    assert blk.shape == (padded, n) and np.count_nonzero(blk[owned:]) == 0
    q = blk
    r_total = None
    for _ in range(2):                                                      # sweep_1d twice (cacqr.hpp:180-183)
        g = torch.from_numpy(orc.syrk_upper_trans(q).copy())
        dist.all_reduce(g)                                                  # the Gram exchange (policy.h:82)
        r = orc.potrf_upper(g.numpy())
        q = orc.trmm_right_upper(q, orc.trtri_upper(r))
        r_total = r if r_total is None else orc.trmm_right_upper(r, r_total)
    ident = cdist.broadcast_bytes(bytes(range(128)) if rank == 0 else b"", 0)
    tmax = cdist.max_over_ranks(10.0 + rank)
    np.savez(os.path.join(out, "r%d.npz" % rank), Q=q, R=r_total, ident=np.frombuffer(ident, dtype=np.uint8), tmax=tmax)
    dist.destroy_process_group()


def test_sharded_cqr2_equals_unsharded(tmp_path):
    import capital_oracle as orc
    m, n, world = 1001, 24, 2                                               # ragged: rank 1 holds one padding row
    mp.spawn(_worker, args=(world, 29517, m, n, str(tmp_path)), nprocs=world, join=True)
    blocks = orc.cacqr_input_blocks(m, n, world)
    q_ref, r_ref = orc.cacqr_factor_1d(blocks, 2)
    for rank in range(world):
        got = np.load(tmp_path / ("r%d.npz" % rank))
        assert np.linalg.norm(got["R"] - r_ref) / np.linalg.norm(r_ref) <= 1e-13
        assert np.linalg.norm(got["Q"] - q_ref[rank]) / np.linalg.norm(q_ref[rank]) <= 1e-13
        assert bytes(got["ident"]) == bytes(range(128))
        assert float(got["tmax"]) == 11.0
    A = orc.assemble_row_cyclic(blocks, m); Q = orc.assemble_row_cyclic(q_ref, m)
    assert orc.qr_orthogonality(Q) <= 1e-13 * n and orc.qr_residual(A, Q, r_ref) <= 1e-13 * n


def test_local_rows():
    from capital_b200.dist import local_rows
    assert local_rows(10, 4, 0) == (3, 3) and local_rows(10, 4, 1) == (3, 3)
    assert local_rows(10, 4, 2) == (3, 2) and local_rows(10, 4, 3) == (3, 2)
    assert local_rows(1 << 22, 8, 5) == (1 << 19, 1 << 19)


def _grid_worker(rank, world, port, c, N, out):
    """The process-grid schedule of csrc/grid_algs.cuh restated with gloo collectives and numpy blocks:
    layer z takes the k-panels kp = z, z+c, ...; left operand from x == kp along the row, right operand from
    y == kp along the column, partial products summed over depth (summa.hpp:164-253 generalised to d % c == 0)."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from capital_b200 import dist as cdist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = int(round((world // c) ** 0.5))
    x, y, z = cdist.grid_coords(rank, c, d)
    ranks = {cdist.grid_coords(r, c, d): r for r in range(world)}
    # every process creates every group, in the same order (torch.distributed requirement)
    row = {(yy, zz): dist.new_group([ranks[(xx, yy, zz)] for xx in range(d)]) for yy in range(d) for zz in range(c)}
    col = {(xx, zz): dist.new_group([ranks[(xx, yy, zz)] for yy in range(d)]) for xx in range(d) for zz in range(c)}
    dep = {(xx, yy): dist.new_group([ranks[(xx, yy, zz)] for zz in range(c)]) for xx in range(d) for yy in range(d)}
    rng = np.random.default_rng(3)
    A, B = rng.random((N, N)), rng.random((N, N))
    a = torch.from_numpy(np.ascontiguousarray(cdist.cyclic_block(A, x, y, d)))
    b = torch.from_numpy(np.ascontiguousarray(cdist.cyclic_block(B, x, y, d)))
    part = torch.zeros_like(a)
    if d == 1:                                   # c x 1 x 1: replicated operands, the k range cut in c slices
        kz = -(-N // c)
        part = a[:, z * kz:(z + 1) * kz] @ b[z * kz:(z + 1) * kz, :]
    else:
        for s in range(d // c):
            kp = z + s * c
            lb = a.clone() if x == kp else torch.empty_like(a)
            dist.broadcast(lb, ranks[(kp, y, z)], group=row[(y, z)])
            rb = b.clone() if y == kp else torch.empty_like(b)
            dist.broadcast(rb, ranks[(x, kp, z)], group=col[(x, z)])
            part += lb @ rb
    if c > 1:
        dist.all_reduce(part, group=dep[(x, y)])
    np.save(os.path.join(out, "c%d.npy" % rank), part.numpy())
    dist.destroy_process_group()


def _check_grid(tmp_path, world, c, port):
    from capital_b200 import dist as cdist
    N = 24
    mp.spawn(_grid_worker, args=(world, port, c, N, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(3)
    A, B = rng.random((N, N)), rng.random((N, N))
    d = int(round((world // c) ** 0.5))
    for zz in range(c):
        blocks = {}
        for r in range(world):
            xx, yy, z = cdist.grid_coords(r, c, d)
            if z == zz:
                blocks[(xx, yy)] = np.load(tmp_path / ("c%d.npy" % r))
        C = cdist.assemble_cyclic(blocks, d)
        assert np.linalg.norm(C - A @ B) / np.linalg.norm(A @ B) <= 1e-14


def test_summa_schedule_2x1x1(tmp_path):
    _check_grid(tmp_path, 2, 2, 29521)


def test_summa_schedule_1x2x2(tmp_path):
    _check_grid(tmp_path, 4, 1, 29522)


def test_summa_schedule_2x2x2(tmp_path):
    _check_grid(tmp_path, 8, 2, 29523)


def test_grid_coords_match_reference_layout0():
    """topo::square layout 0 (reference src/util/topology.h:80-82): z = rank % c, y = rank / (d*c), x = (rank % (d*c)) / c."""
    from capital_b200.dist import default_depth, grid_coords
    assert [grid_coords(r, 2, 2) for r in range(8)] == [(0, 0, 0), (0, 0, 1), (1, 0, 0), (1, 0, 1), (0, 1, 0), (0, 1, 1), (1, 1, 0), (1, 1, 1)]
    assert [grid_coords(r, 1, 2) for r in range(4)] == [(0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 1, 0)]
    assert default_depth(8) == 2 and default_depth(4) == 1 and default_depth(2) == 2 and default_depth(1) == 1
