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
