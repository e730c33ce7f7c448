#!/bin/bash
# 8-GPU: cfg3 (cholinv 32768 on 2x2x2) timings and the synchronised phase profile
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
B=capital_b200/host/bin
$TR --master-port 29602 --no-python $B/cholinv 32768 1 1 1 -2 0 0 3 > gpurun_out/cfg3_ci1.log 2>&1; grep -E "total time|rror" gpurun_out/cfg3_ci1.log | tail -3
$TR --master-port 29603 --no-python $B/cholinv 32768 1 0 1 -2 0 0 3 > gpurun_out/cfg3_ci0.log 2>&1; grep -E "total time|rror" gpurun_out/cfg3_ci0.log | tail -3
CAPITAL_B200_PROFILE=1 $TR --master-port 29604 --no-python $B/cholinv 32768 1 1 1 -2 0 0 2 > gpurun_out/cfg3_ci1_prof.log 2>&1; grep -E "profile|total time|rror" gpurun_out/cfg3_ci1_prof.log | tail -4
CAPITAL_B200_PROFILE=1 $TR --master-port 29605 --no-python $B/cholinv 32768 1 0 1 -2 0 0 2 > gpurun_out/cfg3_ci0_prof.log 2>&1; grep -E "profile|total time|rror" gpurun_out/cfg3_ci0_prof.log | tail -4
