"""Multi-GPU parity check (launched under torchrun, one rank per GPU): sharded CholeskyQR2 through the
C ABI with the NCCL Gram allreduce vs the numpy oracle run over the same rank blocks."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import torch
import torch.distributed as dist

import capital_oracle as orc
from capital_b200 import dist as cdist
from capital_b200.engine import Engine, colmajor

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = Engine(local); comm = cdist.Comm(eng)
m, n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096, int(sys.argv[2]) if len(sys.argv) > 2 else 192
m_loc, owned = cdist.local_rows(m, world, rank)
A = colmajor(m_loc, n, fill=0.0); eng.fill_random(A[:owned], rank)
Q = colmajor(m_loc, n); R = colmajor(n, n)
eng.reset_info(); eng.cacqr_factor_1d(A, Q, R, 2, comm.ptr); torch.cuda.synchronize()
blocks = orc.cacqr_input_blocks(m, n, world)
assert np.array_equal(A.cpu().numpy(), blocks[rank]), "input block differs from the reference generator"
q_ref, r_ref = orc.cacqr_factor_1d(blocks, 2)
rel = lambda x, y: float(np.linalg.norm(x - y) / np.linalg.norm(y))
dq, dr = rel(Q.cpu().numpy(), q_ref[rank]), rel(R.cpu().numpy(), r_ref)
# every rank must hold the same R bit for bit (the allreduce result is replicated)
r_all = [torch.empty_like(R.t().contiguous()) for _ in range(world)]
dist.all_gather(r_all, R.t().contiguous())
same = all(torch.equal(r_all[0], x) for x in r_all)
ok = dq <= 1e-11 and dr <= 1e-11 and same and eng.info() == 0
print("rank %d/%d m=%d n=%d dQ=%.2e dR=%.2e replicated_R=%s %s" % (rank, world, m, n, dq, dr, same, "OK" if ok else "FAIL"), flush=True)
comm.close(); eng.close(); dist.destroy_process_group()
sys.exit(0 if ok else 1)
