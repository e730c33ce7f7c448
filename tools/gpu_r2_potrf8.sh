#!/bin/bash
# Round 2, P GPUs: asynchronous cap_potrf_1d (peer copy-engine exchange, staircase trailing update) against the round-1 NCCL order,
# timelines, then the full bench line.
P=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29631 tools/potrf1d_check.py 4096 256 128 --pipeline -1 --exchange 1 2>&1 | grep -E "rank|capital_b200|rror" | tail -16
timeout 120 $TR --master-port 29632 tools/potrf1d_check.py 4000 128 --pipeline -1 --exchange 1 2>&1 | grep -E "FAIL|capital_b200|rror" | tail -16
timeout 300 $TR --master-port 29622 tools/potrf1d_check.py 32768 1024 512 2048 --time 3 --pipeline -1 0 --reserve 0 16 32 --grid 0 --exchange 1 --json gpurun_out/potrf1d_r02_n$P.jsonl > gpurun_out/potrf1d_multi_a.out 2> gpurun_out/potrf1d_multi_a.err; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_a.out gpurun_out/potrf1d_multi_a.err | tail -30
CAPITAL_B200_STAIR=0 timeout 300 $TR --master-port 29623 tools/potrf1d_check.py 32768 1024 --time 3 --pipeline -1 --reserve 16 --grid 0 --exchange 1 > gpurun_out/potrf1d_multi_b.out 2> gpurun_out/potrf1d_multi_b.err; echo "STAIR=0:"; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_b.out gpurun_out/potrf1d_multi_b.err | tail -30
timeout 120 $TR --master-port 29624 tools/potrf1d_check.py 32768 1024 --pipeline -1 --reserve 16 --exchange 1 --timeline --no-check > gpurun_out/potrf1d_timeline_n$P.out 2> gpurun_out/potrf1d_timeline_n$P.err; grep -c "panel" gpurun_out/potrf1d_timeline_n$P.err
timeout 400 $TR --master-port 29625 bench.py --gpus $P --steps 3 --warmup 3 > gpurun_out/r02_bench_n${P}_a.json 2> gpurun_out/r02_bench_n${P}_a.err; echo bench_rc=$?; tail -c 600 gpurun_out/r02_bench_n${P}_a.err
