#!/bin/bash
# Round 2, P GPUs: asynchronous cap_potrf_1d (peer copy-engine exchange, staircase bulk update beside the row work), per-rank timelines.
P=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29631 tools/potrf1d_check.py 4096 128 --pipeline -1 --exchange 1 2>&1 | grep -E "FAIL|capital_b200|rror| OK" | tail -16
timeout 300 $TR --master-port 29622 tools/potrf1d_check.py 32768 1024 512 2048 --time 3 --pipeline -1 --reserve 0 16 --grid 0 64 --exchange 1 --json gpurun_out/potrf1d_r02b_n$P.jsonl > gpurun_out/potrf1d_multi_a.out 2> gpurun_out/potrf1d_multi_a.err; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_a.out gpurun_out/potrf1d_multi_a.err | tail -30
CAPITAL_B200_PROFILE_FILE=gpurun_out/tl_n$P timeout 120 $TR --master-port 29624 tools/potrf1d_check.py 32768 1024 --pipeline -1 --reserve 16 --grid 64 --exchange 1 --timeline --no-check > /dev/null 2>&1; ls gpurun_out | grep -c tl_n$P
