"""Parameter sweep: the job of the reference's autotune/{cholesky/cholinv,qr/cacqr}/tune.cpp without the external
`critter` profiler -- every configuration is simply run and timed (CUDA events, max over ranks), and the table
config -> seconds / TFLOP/s is printed and written as JSON.  On a B200 the best base-case size differs radically
from the CPU defaults, which is what this is for.

    python tools/sweep.py cholinv --N 16384 --bc 0 -1 -2 -3 -4 --ci 0 1 --split 1
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/sweep.py cholinv --N 32768
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/sweep.py cacqr3d --M 1048576 --N 4096

Knobs (reference names): complete_inv, split, bc_mult_dim (cholinv.hpp:6-18: base case = c*d * 2^-bc_mult_dim local
rows -> global size d * l / that), grid depth c (topo::square), num_iter (1 = CholeskyQR, 2 = CholeskyQR2)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from capital_b200 import dist as cdist
from capital_b200.engine import Engine, colmajor


def bc_global_of(l, c, d, bc_mult_dim):
    """cholinv.hpp:15-18 as mirrored by capital_b200/host/capital.hpp"""
    bc = c * d
    if bc_mult_dim < 0:
        bc <<= -bc_mult_dim
    else:
        bc >>= bc_mult_dim
    bc = min(max(bc, 1), l)
    return d * (l // bc)


def timed(fn, iters, world):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(iters):
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    return best * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("alg", choices=["cholinv", "cacqr3d"])
    ap.add_argument("--N", dest="n", type=int, default=8192)
    ap.add_argument("--M", dest="m", type=int, default=1 << 20)
    ap.add_argument("--bc", type=int, nargs="*", default=[0, -1, -2, -3, -4])
    ap.add_argument("--ci", type=int, nargs="*", default=[0, 1])
    ap.add_argument("--split", type=int, nargs="*", default=[1])
    ap.add_argument("--depth", type=int, nargs="*", default=None, help="grid depths c to try (default: the library's default)")
    ap.add_argument("--num-iter", type=int, default=2)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    eng = Engine(lr)
    comm = cdist.Comm(eng) if world > 1 else None
    rows = []
    for c in (args.depth or [cdist.default_depth(world)]):
        grid = cdist.Grid(eng, comm, c)
        d = grid.d
        l = args.n // d
        if args.alg == "cholinv":
            S = colmajor(l, l); eng.fill_symmetric(S, grid.x, grid.y, d, d, args.n, True)
            W, X = colmajor(l, l), colmajor(l, l)
            for ci in args.ci:
                for split in args.split:
                    for bc in args.bc:
                        bcg = bc_global_of(l, c, d, bc)
                        sec = timed(lambda: (W.copy_(S), eng.cholinv_factor(grid, W, X, bool(ci), split, bcg)), args.iters, world)
                        rows.append({"alg": "cholinv", "N": args.n, "grid": "%dx%dx%d" % (c, d, d), "complete_inv": ci, "split": split,
                                     "bc_mult_dim": bc, "bc_global": bcg, "seconds": sec, "tflops_n3_over_3": args.n ** 3 / 3 / sec * 1e-12})
        else:
            ml = args.m // d
            A = colmajor(ml, l); eng.fill_random(A, 1 + grid.x + d * grid.y)
            Q, R = colmajor(ml, l), colmajor(l, l)
            fl = (4.0 if args.num_iter == 2 else 2.0) * args.m * args.n ** 2 + (5.0 / 3 if args.num_iter == 2 else 2.0 / 3) * args.n ** 3
            for split in args.split:
                for bc in args.bc:
                    bcg = bc_global_of(l, c, d, bc)
                    sec = timed(lambda: eng.cacqr_factor_3d(grid, A, Q, R, args.num_iter, None, split, bcg), args.iters, world)
                    rows.append({"alg": "cacqr3d", "M": args.m, "N": args.n, "grid": "%dx%dx%d" % (c, d, d), "num_iter": args.num_iter, "split": split,
                                 "bc_mult_dim": bc, "bc_global": bcg, "seconds": sec, "tflops": fl / sec * 1e-12})
        grid.close()
    if rank == 0:
        keys = list(rows[0].keys()) if rows else []
        print(" ".join("%-14s" % k for k in keys))
        for r in rows:
            print(" ".join("%-14s" % (("%.5g" % r[k]) if isinstance(r[k], float) else r[k]) for k in keys), flush=True)
        best = min(rows, key=lambda r: r["seconds"]) if rows else None
        print("best:", json.dumps(best))
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w") as f:
            json.dump({"world": world, "rows": rows, "best": best}, f, indent=1)
    if comm:
        comm.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
