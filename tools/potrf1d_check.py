"""cap_potrf_1d (look-ahead Cholesky, block columns dealt 1-D over the GPUs): parity and timing.

    python tools/potrf1d_check.py N nb [--time ITERS] [--reserve SMS]                (one GPU)
    python -m torch.distributed.run --nproc-per-node P ... tools/potrf1d_check.py N nb ...   (P GPUs, one rank each)

Input: the reference's SPD generator (distribute_symmetric, src/matrix/structure.hpp:69-103), bit-exact on device.
Check (every rank, its own block columns): against numpy's Cholesky of the same matrix when N <= 4096, else
against the single-GPU cap_dpotrf of the whole matrix computed redundantly on this rank (recursive schedule, itself
pinned to numpy at small sizes by tests/test_gpu_solve.py).  Tolerance: relative factor difference <= 1e-11.
Timing: CUDA events on the handle's stream around ITERS factorizations (input restored by a device copy that is
timed alone and subtracted), barrier + synchronize on both sides, max over ranks; TFLOP/s on N^3/3.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from capital_b200 import dist as cdist  # noqa: E402
from capital_b200.engine import Engine, colmajor  # noqa: E402

PEAK = 36.6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("N", type=int)
    ap.add_argument("nb", type=int, nargs="+")
    ap.add_argument("--pipeline", type=int, nargs="+", default=[-1], help="-1 = asynchronous schedule (peer copy engines), 0 / 1 = round-1 NCCL orders")
    ap.add_argument("--grid", type=int, nargs="+", default=[0], help="cap_set_update_grid values to time (0 = persistent SM-capped grid)")
    ap.add_argument("--exchange", type=int, default=-1, help="cap_set_peer_exchange: -1 auto, 0 NCCL only, 1 peer required")
    ap.add_argument("--time", type=int, default=0)
    ap.add_argument("--reserve", type=int, nargs="+", default=[16])
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--timeline", action="store_true", help="after the timed runs, one more factorization with CAPITAL_B200_PROFILE=2 "
                    "(per-panel phase timeline of every rank on stderr; warm: no workspace growth inside)")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    eng.set_peer_exchange(args.exchange)
    comm = cdist.Comm(eng) if world > 1 else None
    cptr = comm.ptr if comm else None
    N = args.N
    all_ok = True
    results = []
    ref_cache = {}
    for nb in args.nb:
        eng.set_lookahead(args.reserve[0], args.pipeline[0])
        mine = cdist.potrf_1d_blocks(N, nb, world, rank)
        widths = [min(nb, N - j * nb) for j in mine]
        ncols = sum(widths)
        A0 = colmajor(N, max(ncols, 1))
        for r, (j, w) in enumerate(zip(mine, widths)):      # global columns [j nb, j nb + w) of the SPD matrix
            eng.fill_symmetric(A0[:, r * nb:r * nb + w], j * nb, 0, 1, 1, N, True)
            if j > 0:   # the generator adds N only where LOCAL row == LOCAL column (structure.hpp:87-89): do it for the shifted block
                A0[:, r * nb:r * nb + w].diagonal(-j * nb).add_(float(N))
        A = colmajor(N, max(ncols, 1)); A.copy_(A0)
        eng.reset_info()
        eng.potrf_1d(A, N, nb, cptr)
        torch.cuda.synchronize()
        def check():
            torch.cuda.synchronize()
            good, worst = eng.info() == 0, 0.0
            if args.no_check:
                return good, worst
            if N <= 4096:
                import capital_oracle as orc
                if "np" not in ref_cache:
                    ref_cache["np"] = np.linalg.cholesky(orc.symmetric_global(N)).T
                get = lambda j, w: torch.from_numpy(np.ascontiguousarray(ref_cache["np"][:, j * nb:j * nb + w])).cuda()
            else:
                if "F" not in ref_cache:     # the whole matrix factored redundantly on this rank (recursive single-GPU schedule)
                    F = colmajor(N, N); eng.fill_symmetric(F, 0, 0, 1, 1, N, True)
                    eng.set_potrf_schedule(1); eng._potrf(F, N, "U"); torch.cuda.synchronize()
                    ref_cache["F"] = F
                get = lambda j, w: ref_cache["F"][:, j * nb:j * nb + w]
            for r, (j, w) in enumerate(zip(mine, widths)):
                got, want = torch.triu(A[:, r * nb:r * nb + w], -j * nb), torch.triu(get(j, w), -j * nb)
                worst = max(worst, float((got - want).norm() / want.norm()))
                low_got, low_in = torch.tril(A[:, r * nb:r * nb + w], -j * nb - 1), torch.tril(A0[:, r * nb:r * nb + w], -j * nb - 1)
                good = good and bool(torch.equal(low_got, low_in))               # nothing below the diagonal is written
            return good and worst <= 1e-11, worst
        ok, worst = check()
        print("rank %d/%d N=%d nb=%d blocks=%s rel=%.2e info=%d %s" % (rank, world, N, nb, mine if len(mine) <= 8 else len(mine), worst, eng.info(),
                                                                       "OK" if ok else "FAIL"), flush=True)
        all_ok = all_ok and ok
        if args.time > 0:
            def barrier():
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()

            def timed(fn):
                barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(args.time):
                    fn()
                b.record(); barrier()
                return cdist.max_over_ranks(a.elapsed_time(b)) / args.time
            t_copy = timed(lambda: A.copy_(A0))
            for pipeline in args.pipeline:
                for reserve in args.reserve:
                  for grid in args.grid:
                    eng.set_lookahead(reserve, pipeline); eng.set_update_grid(grid)
                    for _ in range(2):
                        A.copy_(A0); eng.potrf_1d(A, N, nb, cptr)
                    t = timed(lambda: (A.copy_(A0), eng.potrf_1d(A, N, nb, cptr))) - t_copy
                    tf = N ** 3 / 3 / t * 1e-9
                    good, worst = check()
                    all_ok = all_ok and good
                    results.append({"N": N, "nb": nb, "gpus": world, "reserve_sms": reserve, "pipeline": pipeline, "update_grid": grid, "ms": t, "tflops": tf,
                                    "frac_of_peak": tf / (PEAK * world), "rel_vs_reference": worst, "ok": bool(good)})
                    if rank == 0:
                        print("potrf_1d N=%d nb=%d on %d GPU(s), pipeline %d, %d SMs reserved, grid x%d: %.3f ms  %.2f TFLOP/s  %.1f %% of %d x %.1f  (rank 0: rel %.1e %s)" % (
                            N, nb, world, pipeline, reserve, grid, t, tf, 100 * tf / (PEAK * world), world, PEAK, worst, "OK" if good else "FAIL"), flush=True)
        if args.timeline:
            for _ in range(2):
                A.copy_(A0); eng.potrf_1d(A, N, nb, cptr)
            torch.cuda.synchronize()
            os.environ["CAPITAL_B200_PROFILE"] = "2"
            A.copy_(A0); eng.potrf_1d(A, N, nb, cptr); torch.cuda.synchronize()
            os.environ.pop("CAPITAL_B200_PROFILE", None)
        del A, A0
        torch.cuda.empty_cache()
    if rank == 0 and args.json and results:
        with open(args.json, "a") as f:
            for r in results:
                f.write(json.dumps(r) + "\n")
    eng.close()          # the handle owns a duplicate of comm: it goes first
    if comm:
        comm.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if all_ok else 1)


if __name__ == "__main__":
    main()
