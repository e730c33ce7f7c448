#!/bin/bash
# Round 2 final, P GPUs: cap_potrf_1d block-width sweep with per-rank timelines, then the full bench line.
P=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29622 tools/potrf1d_check.py 32768 1024 768 1536 512 --time 3 --pipeline -1 --reserve 16 --grid 64 --exchange 1 --json gpurun_out/potrf1d_r02c_n$P.jsonl > gpurun_out/potrf1d_multi_a.out 2> gpurun_out/potrf1d_multi_a.err; grep -E "^potrf_1d|FAIL|rror|capital_b200" gpurun_out/potrf1d_multi_a.out gpurun_out/potrf1d_multi_a.err | tail -30
CAPITAL_B200_PROFILE_FILE=gpurun_out/tlc_n$P timeout 120 $TR --master-port 29624 tools/potrf1d_check.py 32768 1024 --pipeline -1 --reserve 16 --grid 64 --exchange 1 --timeline --no-check > /dev/null 2>&1; ls gpurun_out | grep -c tlc_n$P
timeout 500 $TR --master-port 29625 bench.py --gpus $P --steps 5 --warmup 3 > gpurun_out/r02_bench_n${P}_b.json 2> gpurun_out/r02_bench_n${P}_b.err; echo bench_rc=$?; tail -c 300 gpurun_out/r02_bench_n${P}_b.err
