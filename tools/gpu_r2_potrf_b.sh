#!/bin/bash
P=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
echo "== pytest (potrf tests)"; timeout 600 python -m pytest tests/test_gpu_solve.py -x -q -m gpu -k "potrf" 2>&1 | tail -4
timeout 200 python tools/potrf1d_check.py 16384 1024 --time 3 --pipeline -1 --reserve 0 16 --grid 0 64 --json gpurun_out/potrf1d_r02b_n1.jsonl 2>&1 | grep -E "^potrf_1d|FAIL|rror"
timeout 200 python tools/potrf1d_check.py 16384 2048 --time 3 --pipeline -1 --reserve 0 16 --grid 0 64 --json gpurun_out/potrf1d_r02b_n1.jsonl 2>&1 | grep -E "^potrf_1d|FAIL|rror"
timeout 200 python tools/potrf1d_check.py 32768 2048 --time 2 --pipeline -1 --reserve 0 16 --grid 0 64 --json gpurun_out/potrf1d_r02b_n1.jsonl 2>&1 | grep -E "^potrf_1d|FAIL|rror"
CAPITAL_B200_PROFILE_FILE=gpurun_out/tl_n1 timeout 200 python tools/potrf1d_check.py 16384 1024 --pipeline -1 --reserve 0 --grid 64 --timeline --no-check > /dev/null 2>&1
echo "== $P GPUs"
timeout 200 $TR --master-port 29631 tools/potrf1d_check.py 4000 128 256 --pipeline -1 --exchange 1 2>&1 | grep -E "FAIL|capital_b200|rror| OK" | tail -8
timeout 300 $TR --master-port 29622 tools/potrf1d_check.py 32768 1024 2048 --time 3 --pipeline -1 --reserve 0 16 --grid 0 64 --exchange 1 --json gpurun_out/potrf1d_r02b_n$P.jsonl > gpurun_out/potrf1d_multi_a.out 2> gpurun_out/potrf1d_multi_a.err; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_a.out gpurun_out/potrf1d_multi_a.err | tail
CAPITAL_B200_PROFILE_FILE=gpurun_out/tl_n$P timeout 200 $TR --master-port 29624 tools/potrf1d_check.py 32768 1024 --pipeline -1 --reserve 0 --grid 64 --exchange 1 --timeline --no-check > /dev/null 2>&1; ls gpurun_out | grep tl_
