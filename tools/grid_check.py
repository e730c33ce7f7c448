"""Multi-GPU parity of the process-grid algorithms (SUMMA gemm/trmm/syrk, distributed cholinv, 3D CholeskyQR) against
numpy on the assembled global matrices (capital_b200.checks.grid_small_check).
torchrun --nproc-per-node P tools/grid_check.py [N]   (P in 1, 2, 4, 8)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from capital_b200 import checks
from capital_b200 import dist as cdist
from capital_b200.engine import Engine


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    eng = Engine(lr)
    comm = cdist.Comm(eng) if world > 1 else None
    res = checks.grid_small_check(eng, comm, world, rank, N, verbose=True, reference_fixture=True)
    if rank == 0:
        print("GRID CHECK", "PASSED" if res["ok"] else "FAILED", "world", world, "grid", res["grid"], "N", N, flush=True)
    eng.close()
    if comm:
        comm.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
