"""One process per GPU: torch.distributed is the plumbing (rendezvous, id exchange, timing reduction);
the data-path collectives are NCCL calls inside libcapital_b200.so (csrc/comm.cu).

Sharding of the CholeskyQR path follows the reference's 1D layout (bench/qr/cacqr.cpp:32-34):
matrix A(n, m, c=1, d=P) -- rows dealt cyclically, rank r owns ceil-divided rows and generates them
with distribute_random(key = rank)."""
import ctypes as C

import torch
import torch.distributed as dist


def local_rows(m, ranks, rank):
    """Rows of the m x n matrix owned by `rank` under the reference's cyclic deal; the local buffer has
    ceil(m/P) rows (matrix.hpp:8-11), of which the last is zero padding on ranks past the remainder."""
    padded = -(-m // ranks)
    owned = padded if (m % ranks == 0 or rank < m % ranks) else padded - 1
    return padded, owned


def broadcast_bytes(payload, src=0):
    """Broadcast a bytes object from `src` over the default process group (gloo or nccl)."""
    n = torch.tensor([len(payload) if dist.get_rank() == src else 0], dtype=torch.int64)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    n = n.to(dev); dist.broadcast(n, src)
    buf = torch.zeros(int(n.item()), dtype=torch.uint8, device=dev)
    if dist.get_rank() == src:
        buf.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(buf, src)
    return bytes(buf.cpu().numpy().tobytes())


def max_over_ranks(value):
    """Device-timed milliseconds -> max over ranks (the reference's PMPI_Allreduce(MAX), bench/qr/cacqr.cpp:52)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def potrf_1d_blocks(n, nb, ranks, rank):
    """Global block-column indices held by `rank` under cap_potrf_1d's boustrophedon deal, in local order:
    block j belongs to round r = j // P and to rank (j % P if r is even else P-1 - j % P), where it is local block r."""
    nblk = -(-n // nb)
    out = []
    r = 0
    while True:
        j = r * ranks + (ranks - 1 - rank if r & 1 else rank)
        if j >= nblk:
            return out
        out.append(j)
        r += 1


class Comm:
    """NCCL communicator owned by the C-ABI library (cap_comm_t), bootstrapped through torch.distributed."""

    def __init__(self, engine):
        self.lib = engine.lib
        self.ptr = C.c_void_p()
        self.rank, self.size = dist.get_rank(), dist.get_world_size()
        ident = C.create_string_buffer(128)
        if self.rank == 0:
            from . import _lib
            _lib.check(self.lib.cap_comm_unique_id(ident), "cap_comm_unique_id")
        raw = broadcast_bytes(ident.raw, 0)
        from . import _lib
        _lib.check(self.lib.cap_comm_init_rank(C.byref(self.ptr), raw, self.size, self.rank), "cap_comm_init_rank")

    def split(self, color, key):
        from . import _lib
        child = Comm.__new__(Comm)
        child.lib, child.ptr = self.lib, C.c_void_p()
        _lib.check(self.lib.cap_comm_split(self.ptr, color, key, C.byref(child.ptr)), "cap_comm_split")
        r, s = C.c_int(), C.c_int()
        self.lib.cap_comm_rank(child.ptr, C.byref(r)); self.lib.cap_comm_size(child.ptr, C.byref(s))
        child.rank, child.size = r.value, s.value
        return child

    def close(self):
        if self.ptr:
            self.lib.cap_comm_destroy(self.ptr)
            self.ptr = C.c_void_p()


def default_depth(world):
    """Depth c of the c x d x d grid used for `world` GPUs: 1 -> 1x1x1, 2 -> 2x1x1 (replicated operands, k split),
    4 -> 1x2x2 (2D SUMMA), 8 -> 2x2x2 (the reference's cube, topo::square needs P = c^3)."""
    return {1: 1, 2: 2, 4: 1, 8: 2}[world]


def grid_coords(rank, c, d):
    """rank -> (x, y, z), layout 0 of topo::square (reference src/util/topology.h:80-82)."""
    return (rank % (d * c)) // c, rank // (d * c), rank % c


def cyclic_block(a, x, y, d):
    """Local block of a global array under the element-cyclic deal (structure.hpp:74-98): rows y::d, columns x::d."""
    return a[y::d, x::d]


def assemble_cyclic(blocks, d):
    """Inverse of cyclic_block: blocks[(x, y)] -> global array."""
    r, c = blocks[(0, 0)].shape
    out = blocks[(0, 0)].new_empty((r * d, c * d)) if hasattr(blocks[(0, 0)], "new_empty") else None
    if out is None:
        import numpy as np
        out = np.empty((r * d, c * d), dtype=blocks[(0, 0)].dtype)
    for (x, y), b in blocks.items():
        out[y::d, x::d] = b
    return out


class Grid:
    """topo::square over the library's NCCL communicator (cap_grid_t)."""

    def __init__(self, engine, comm, c, layout=0, num_chunks=0):
        """layout 0 / 1: the reference's two rank orders (only layout 0 matches grid_coords); num_chunks > 0 selects the
        pipelined SUMMA panel loop (summa.hpp:177-215)."""
        from . import _lib
        self.lib = engine.lib
        self.ptr = C.c_void_p()
        _lib.check(self.lib.cap_grid_create2(comm.ptr if comm else None, int(c), int(layout), int(num_chunks), C.byref(self.ptr)), "cap_grid_create")
        v = [C.c_int() for _ in range(5)]
        self.lib.cap_grid_info(self.ptr, *[C.byref(t) for t in v])
        self.c, self.d, self.x, self.y, self.z = [t.value for t in v]

    def close(self):
        if self.ptr:
            self.lib.cap_grid_destroy(self.ptr)
            self.ptr = C.c_void_p()
