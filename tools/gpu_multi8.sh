#!/bin/bash
# 8-GPU session: headline bench at N=8, cfg3 (cholinv 32768 on 2x2x2), cfg4 (SUMMA 65536 on 2x2x2)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29601 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
tail -c 2500 gpurun_out/bench_n8.json; tail -3 gpurun_out/bench_n8.err
B=capital_b200/host/bin
$TR --master-port 29602 --no-python $B/cholinv 32768 1 1 1 -2 0 0 2 > gpurun_out/cfg3_ci1.log 2>&1; grep -E "total time|error|Error" gpurun_out/cfg3_ci1.log | tail -4
$TR --master-port 29603 --no-python $B/cholinv 32768 1 0 1 -2 0 0 2 > gpurun_out/cfg3_ci0.log 2>&1; grep -E "total time|error|Error" gpurun_out/cfg3_ci0.log | tail -4
$TR --master-port 29604 --no-python $B/cholinv 32768 1 0 1 -3 0 0 2 > gpurun_out/cfg3_ci0_bc3.log 2>&1; grep -E "total time|error|Error" gpurun_out/cfg3_ci0_bc3.log | tail -4
$TR --master-port 29605 --no-python $B/summa_gemm 65536 65536 65536 1 0 0 2 > gpurun_out/cfg4.log 2>&1; grep -E "TFLOP|error|Error" gpurun_out/cfg4.log | tail -4
