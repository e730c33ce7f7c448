"""Host-side logic of cap_potrf_1d (csrc/potrf_1d.cuh) on CPU: the boustrophedon block-column deal exported by the
C ABI agrees with the Python helper, balances the triangular work, and a numpy walk-through of the schedule that
uses exactly the index formulas of the CUDA host code (owner / upto / cols_right / run detection / the gathered-panel
ordering) reproduces LAPACK's Cholesky factor for every rank count and ragged size."""
import ctypes as C

import numpy as np
import pytest

from capital_b200 import _lib, dist as cdist


def c_layout(n, nb, ranks, rank):
    lib = _lib.load()
    cols, blocks = C.c_int64(), C.c_int()
    _lib.check(lib.cap_potrf_1d_layout(n, nb, ranks, rank, C.byref(cols), C.byref(blocks)), "layout")
    return cols.value, blocks.value


def c_block(j, ranks):
    lib = _lib.load()
    own, loc = C.c_int(), C.c_int()
    _lib.check(lib.cap_potrf_1d_block(j, ranks, C.byref(own), C.byref(loc)), "block")
    return own.value, loc.value


@pytest.mark.parametrize("n,nb,ranks", [(1, 128, 1), (1000, 128, 3), (4096, 256, 8), (32768, 1024, 8), (5000, 512, 4), (300, 128, 8)])
def test_layout_matches_python_helper(n, nb, ranks):
    nblk = -(-n // nb)
    seen = {}
    for p in range(ranks):
        mine = cdist.potrf_1d_blocks(n, nb, ranks, p)
        cols, blocks = c_layout(n, nb, ranks, p)
        assert blocks == len(mine)
        assert cols == sum(min(nb, n - j * nb) for j in mine)
        assert mine == sorted(mine)                                          # local order = global order
        for r, j in enumerate(mine):
            assert c_block(j, ranks) == (p, r)
            seen[j] = p
    assert sorted(seen) == list(range(nblk))                                 # every block has exactly one owner


def test_layout_balances_the_upper_triangle():
    """Work of block column j in a right-looking upper Cholesky ~ (j+1)^2 / 2 block products (summed over panels)."""
    n, nb, ranks = 32768, 512, 8
    work = [sum((j + 1) ** 2 for j in cdist.potrf_1d_blocks(n, nb, ranks, p)) for p in range(ranks)]
    assert max(work) / (sum(work) / ranks) < 1.06
    plain = [sum((j + 1) ** 2 for j in range(p, n // nb, ranks)) for p in range(ranks)]
    assert max(plain) / (sum(plain) / ranks) > 1.15                          # what the plain cyclic deal would cost


class Layout:
    """The formulas of struct Layout1D, verbatim."""

    def __init__(self, P, nb, N):
        self.P, self.nb, self.N, self.nblk = P, nb, N, -(-N // nb)

    def owner(self, j):
        r, pos = divmod(j, self.P)
        return self.P - 1 - pos if r & 1 else pos

    def global_block(self, r, p):
        return r * self.P + (self.P - 1 - p if r & 1 else p)

    def width(self, j):
        return min(self.nb, self.N - j * self.nb)

    def nloc(self, p):
        n = 0
        while self.global_block(n, p) < self.nblk:
            n += 1
        return n

    def upto(self, p, j):
        rj, pos = divmod(j, self.P)
        return rj + (1 if (self.P - 1 - p if rj & 1 else p) <= pos else 0)

    def cols(self, p):
        n = self.nloc(p)
        return 0 if n == 0 else (n - 1) * self.nb + self.width(self.global_block(n - 1, p))

    def cols_right(self, p, j):
        return max(0, self.cols(p) - min(self.upto(p, j), self.nloc(p)) * self.nb)


def walk(a, nb, P):
    """numpy walk-through of Potrf1D::run over P virtual ranks; returns the assembled upper factor."""
    N = a.shape[0]
    L = Layout(P, nb, N)
    loc = [np.concatenate([a[:, j * nb:j * nb + L.width(j)] for j in cdist.potrf_1d_blocks(N, nb, P, p)] or [np.zeros((N, 0))], axis=1).copy()
           for p in range(P)]
    S = [None] * P
    Rrow = None

    def panel(j):
        nonlocal Rrow
        own, w = L.owner(j), L.width(j)
        blk = loc[own][j * nb:j * nb + w, (j // P) * nb:(j // P) * nb + w]
        r = np.linalg.cholesky(np.triu(blk) + np.triu(blk, 1).T).T
        blk[np.triu_indices(w)] = r[np.triu_indices(w)]
        if j == L.nblk - 1:
            return
        x = np.linalg.inv(r)                                                  # broadcast
        maxcl = max(L.cols_right(p, j) for p in range(P))
        gath = np.full((P, nb * maxcl), np.nan)
        for p in range(P):
            r0, ncl = min(L.upto(p, j), L.nloc(p)), L.cols_right(p, j)
            if ncl > 0:
                bp = loc[p][j * nb:(j + 1) * nb, r0 * nb:r0 * nb + ncl]
                bp[...] = x.T @ bp
                S[p] = bp.copy()
                gath[p, :nb * ncl] = S[p].T.reshape(-1)                       # column-major nb x ncl
            else:
                S[p] = np.zeros((nb, 0))
        # assemble_row_panel_kernel: block l > j comes from rank owner(l), local position l // P - upto(owner, j)
        parts = []
        for l in range(j + 1, L.nblk):
            own_l, w_l = L.owner(l), L.width(l)
            off = (l // P - L.upto(own_l, j)) * nb * nb
            parts.append(gath[own_l, off:off + nb * w_l].reshape(w_l, nb).T)
        Rrow = np.concatenate(parts, axis=1)
        assert not np.isnan(Rrow).any()

    def update_next_row(j, p):
        k, wk = j + 1, L.width(j + 1)
        r0 = min(L.upto(p, j), L.nloc(p))
        rr = r0
        if p == L.owner(k):
            sk = S[p][:, (k // P - r0) * nb:(k // P - r0) * nb + wk]
            d = loc[p][k * nb:k * nb + wk, (k // P) * nb:(k // P) * nb + wk]
            d[np.triu_indices(wk)] -= (sk.T @ sk)[np.triu_indices(wk)]
            rr = k // P + 1
        ncl = L.cols(p) - min(rr, L.nloc(p)) * nb
        if ncl > 0:
            loc[p][k * nb:k * nb + wk, rr * nb:rr * nb + ncl] -= Rrow[:, :wk].T @ S[p][:, (rr - r0) * nb:(rr - r0) * nb + ncl]

    def update_rest(j, p):
        n = L.nloc(p)
        r0 = min(L.upto(p, j), n)
        r = r0
        while r < n and L.global_block(r, p) < j + 2:
            r += 1
        while r < n:
            r_end = r + 1
            while r_end < n and L.global_block(r_end, p) == L.global_block(r_end - 1, p) + 1:
                r_end += 1
            k0 = L.global_block(r, p)
            wrun = (r_end - 1 - r) * nb + L.width(L.global_block(r_end - 1, p))
            srun = S[p][:, (r - r0) * nb:(r - r0) * nb + wrun]
            hrect = (k0 - j - 2) * nb
            if hrect > 0:
                loc[p][(j + 2) * nb:(j + 2) * nb + hrect, r * nb:r * nb + wrun] -= Rrow[:, nb:nb + hrect].T @ srun
            sq = loc[p][k0 * nb:k0 * nb + wrun, r * nb:r * nb + wrun]
            sq[np.triu_indices(wrun)] -= (srun.T @ srun)[np.triu_indices(wrun)]
            r = r_end

    panel(0)
    for j in range(L.nblk - 1):
        s_prev, r_prev = [s.copy() for s in S], Rrow.copy()                   # panel(j+1) must not disturb U2(j)'s operands
        for p in range(P):
            update_next_row(j, p)
        s_j, r_j = S[:], Rrow
        panel(j + 1)
        s_next, r_next = S[:], Rrow
        if j + 2 < L.nblk:
            S[:], Rrow = s_prev, r_prev
            for p in range(P):
                update_rest(j, p)
            S[:], Rrow = s_next, r_next
        del s_j, r_j
    out = np.zeros((N, N))
    for p in range(P):
        for r, j in enumerate(cdist.potrf_1d_blocks(N, nb, P, p)):
            out[:, j * nb:j * nb + L.width(j)] = loc[p][:, r * nb:r * nb + L.width(j)]
    return out


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("n,nb", [(8, 8), (40, 8), (67, 8), (130, 8), (96, 16)])
def test_schedule_walkthrough_matches_lapack(P, n, nb):
    rng = np.random.default_rng(n + P)
    m = rng.random((n, n)) - 0.5
    a = m @ m.T + n * np.eye(n)
    out = walk(a.copy(), nb, P)
    r = np.linalg.cholesky(a).T
    assert np.allclose(np.triu(out), r, rtol=1e-11, atol=1e-12)
    assert np.array_equal(np.tril(out, -1), np.tril(a, -1))                  # nothing below the diagonal is written


def _gloo_worker(rank, world, port, n, nb, out):
    """cap_potrf_1d's communication pattern with real collectives over gloo: broadcast of the inverse diagonal block from
    the owner, all-gather of equally sized (padded) row-panel pieces, ordering by the closed-form owner / position
    formulas -- one process per rank, each holding only its own block columns."""
    import os
    import sys
    from conftest import ROOT
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(n)
    m = rng.random((n, n)) - 0.5
    a = m @ m.T + n * np.eye(n)
    L = Layout(world, nb, n)
    mine = cdist.potrf_1d_blocks(n, nb, world, rank)
    loc = np.concatenate([a[:, j * nb:j * nb + L.width(j)] for j in mine] or [np.zeros((n, 0))], axis=1).copy()
    for j in range(L.nblk):
        own, w = L.owner(j), L.width(j)
        x = torch.zeros((nb, nb), dtype=torch.float64)
        if rank == own:
            blk = loc[j * nb:j * nb + w, (j // world) * nb:(j // world) * nb + w]
            r = np.linalg.cholesky(np.triu(blk) + np.triu(blk, 1).T).T
            blk[np.triu_indices(w)] = r[np.triu_indices(w)]
            x[:w, :w] = torch.from_numpy(np.linalg.inv(r))
        if j == L.nblk - 1:
            break
        dist.broadcast(x, own)                                                # cap_bcast of X_j
        r0, ncl = min(L.upto(rank, j), L.nloc(rank)), L.cols_right(rank, j)
        maxcl = max(L.cols_right(p, j) for p in range(world))
        send = torch.full((nb * maxcl,), float("nan"), dtype=torch.float64)     # padding is never read
        s = np.zeros((nb, 0))
        if ncl > 0:
            bp = loc[j * nb:(j + 1) * nb, r0 * nb:r0 * nb + ncl]
            bp[...] = x.numpy().T @ bp
            s = bp.copy()
            send[:nb * ncl] = torch.from_numpy(s.T.reshape(-1).copy())
        gath = [torch.empty_like(send) for _ in range(world)]
        dist.all_gather(gath, send)                                            # cap_allgather of the row panel
        parts = []
        for l in range(j + 1, L.nblk):
            own_l, w_l = L.owner(l), L.width(l)
            off = (l // world - L.upto(own_l, j)) * nb * nb
            parts.append(gath[own_l][off:off + nb * w_l].numpy().reshape(w_l, nb).T)
        rrow = np.concatenate(parts, axis=1)
        assert not np.isnan(rrow).any()
        for k_loc in range(r0, L.nloc(rank)):                                  # trailing update of the own block columns
            k = L.global_block(k_loc, rank)
            wk = L.width(k)
            sk = s[:, (k_loc - r0) * nb:(k_loc - r0) * nb + wk]
            rows = rrow[:, :(k - j - 1) * nb + wk]                             # R_j, (j+1 .. k)
            upd = rows.T @ sk
            tgt = loc[(j + 1) * nb:(j + 1) * nb + upd.shape[0], k_loc * nb:k_loc * nb + wk]
            keep = np.triu(np.ones_like(upd), -(k - j - 1) * nb) > 0           # global row <= global column
            tgt[keep] -= upd[keep]
    np.save(os.path.join(out, "p%d.npy" % rank), loc)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,nb,port", [(2, 70, 8, 29581), (3, 100, 16, 29582)])
def test_schedule_over_gloo_matches_lapack(tmp_path, world, n, nb, port):
    import torch.multiprocessing as mp
    mp.spawn(_gloo_worker, args=(world, port, n, nb, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(n)
    m = rng.random((n, n)) - 0.5
    a = m @ m.T + n * np.eye(n)
    r = np.linalg.cholesky(a).T
    L = Layout(world, nb, n)
    for p in range(world):
        loc = np.load(tmp_path / ("p%d.npy" % p))
        for k_loc, j in enumerate(cdist.potrf_1d_blocks(n, nb, world, p)):
            got = loc[:, k_loc * nb:k_loc * nb + L.width(j)]
            want = r[:, j * nb:j * nb + L.width(j)]
            rows = (j + 1) * nb
            assert np.allclose(np.triu(got[:rows], -j * nb), np.triu(want[:rows], -j * nb), rtol=1e-11, atol=1e-12)
            assert np.array_equal(np.tril(got, -j * nb - 1), np.tril(a[:, j * nb:j * nb + L.width(j)], -j * nb - 1))


@pytest.mark.parametrize("P,nb,nblk,ragged", [(1, 128, 9, 0), (1, 512, 7, 100), (2, 256, 11, 0), (3, 384, 10, 200), (4, 128, 17, 1), (8, 1024, 32, 0), (8, 256, 37, 129)])
def test_staircase_enumeration_covers_the_trailing_update_exactly(P, nb, nblk, ragged):
    """One launch = the whole trailing update of a rank (SCHED_STAIR, dmma_tma_gemm.cuh): the kernel's own tile decoder,
    called on the host through cap_debug_stair_decode, must visit every 128 x 128 tile of the rank's block columns >= j+3 that
    lies on or above the GLOBAL diagonal exactly once, rows counted from block row j+2, and flag exactly the diagonal tiles."""
    lib = _lib.load()
    T = nb // 128
    N = nblk * nb - (nb - ragged if ragged else 0)
    L = Layout(P, nb, N)
    for j in range(0, max(1, L.nblk - 2), max(1, L.nblk // 5)):
        for me in range(P):
            mine = cdist.potrf_1d_blocks(N, nb, P, me)
            r = next((i for i, g in enumerate(mine) if g >= j + 3), None)
            if r is None:
                continue
            cols = sum(min(nb, N - g * nb) for g in mine[r:])
            tiles_n = -(-cols // 128)
            row0 = j + 2
            want = {}
            for bi, g in enumerate(mine[r:]):
                for ct in range(-(-min(nb, N - g * nb) // 128)):
                    for tm in range((g - row0) * T + ct + 1):
                        want[(tm, bi * T + ct)] = (tm == (g - row0) * T + ct)
            tm, tn, tri = C.c_int(), C.c_int(), C.c_int()
            units = lib.cap_debug_stair_decode(T, P, me, r, row0, tiles_n, -1, None, None, None)
            assert units == len(want), (j, me, units, len(want))
            got = {}
            for u in range(units):
                lib.cap_debug_stair_decode(T, P, me, r, row0, tiles_n, u, C.byref(tm), C.byref(tn), C.byref(tri))
                assert (tm.value, tn.value) not in got
                got[(tm.value, tn.value)] = bool(tri.value)
            assert got == want
