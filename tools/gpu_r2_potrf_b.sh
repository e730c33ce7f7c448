#!/bin/bash
P=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29631 tools/potrf1d_check.py 4096 256 128 --pipeline -1 --exchange 1 2>&1 | grep -E "rank|capital_b200|rror" | tail -8
timeout 200 $TR --master-port 29632 tools/potrf1d_check.py 4000 128 --pipeline -1 --exchange 1 2>&1 | grep -E "rank|capital_b200|rror" | tail -8
timeout 300 $TR --master-port 29622 tools/potrf1d_check.py 32768 1024 2048 --time 3 --pipeline -1 0 --reserve 16 --grid 0 --exchange 1 --json gpurun_out/potrf1d_r02_n$P.jsonl > gpurun_out/potrf1d_multi_a.out 2> gpurun_out/potrf1d_multi_a.err; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_a.out gpurun_out/potrf1d_multi_a.err | tail
timeout 300 $TR --master-port 29623 tools/potrf1d_check.py 32768 1024 --time 3 --pipeline -1 --reserve 0 --grid 1 2 --exchange 1 --json gpurun_out/potrf1d_r02_n$P.jsonl > gpurun_out/potrf1d_multi_b.out 2> gpurun_out/potrf1d_multi_b.err; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_b.out gpurun_out/potrf1d_multi_b.err | tail
timeout 200 $TR --master-port 29624 tools/potrf1d_check.py 32768 1024 --pipeline -1 --exchange 1 --timeline --no-check > gpurun_out/potrf1d_timeline_n$P.out 2> gpurun_out/potrf1d_timeline_n$P.err; grep -c "panel" gpurun_out/potrf1d_timeline_n$P.err
