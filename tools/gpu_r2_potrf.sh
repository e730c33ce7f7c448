#!/bin/bash
# Round 2: the asynchronous cap_potrf_1d on 1 and P GPUs -- parity first, then timing against the round-1 orders.
#   gpurun --gpus P --timeout 900 -- 'bash tools/gpu_r2_potrf.sh P'
P=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
echo "== pytest (potrf tests)"; timeout 600 python -m pytest tests/test_gpu_solve.py -x -q -m gpu -k "potrf" 2>&1 | tail -5
echo "== 1 GPU timing"
timeout 200 python tools/potrf1d_check.py 16384 1024 2048 --time 3 --pipeline -1 1 --reserve 16 --grid 0 --json gpurun_out/potrf1d_r02_n1.jsonl 2>&1 | grep -E "^potrf_1d|FAIL|rror"
timeout 200 python tools/potrf1d_check.py 16384 1024 --time 3 --pipeline -1 --reserve 0 --grid 1 2 4 --json gpurun_out/potrf1d_r02_n1.jsonl 2>&1 | grep -E "^potrf_1d|FAIL|rror"
timeout 200 python tools/potrf1d_check.py 32768 2048 1024 --time 2 --pipeline -1 --reserve 16 --grid 0 --json gpurun_out/potrf1d_r02_n1.jsonl 2>&1 | grep -E "^potrf_1d|FAIL|rror"
echo "== $P GPU parity + timing"
timeout 200 $TR --master-port 29621 tools/potrf1d_check.py 4096 256 --pipeline -1 --exchange 1 2>&1 | grep -E "rank|FAIL|rror" | tail -12
timeout 300 $TR --master-port 29622 tools/potrf1d_check.py 32768 1024 2048 --time 3 --pipeline -1 0 --reserve 16 --grid 0 --json gpurun_out/potrf1d_r02_n$P.jsonl > gpurun_out/potrf1d_multi_a.out 2> gpurun_out/potrf1d_multi_a.err; grep -E "^potrf_1d|FAIL|rror" gpurun_out/potrf1d_multi_a.out gpurun_out/potrf1d_multi_a.err | tail
timeout 300 $TR --master-port 29623 tools/potrf1d_check.py 32768 1024 --time 3 --pipeline -1 --reserve 0 --grid 2 4 --json gpurun_out/potrf1d_r02_n$P.jsonl > gpurun_out/potrf1d_multi_b.out 2> gpurun_out/potrf1d_multi_b.err; grep -E "^potrf_1d|FAIL|rror" gpurun_out/potrf1d_multi_b.out gpurun_out/potrf1d_multi_b.err | tail
timeout 200 $TR --master-port 29624 tools/potrf1d_check.py 32768 1024 --pipeline -1 --timeline --no-check > gpurun_out/potrf1d_timeline_n$P.out 2> gpurun_out/potrf1d_timeline_n$P.err; grep -c "panel" gpurun_out/potrf1d_timeline_n$P.err
