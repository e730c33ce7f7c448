#!/bin/bash
# 8-GPU session: headline bench at N=8, cfg3 (cholinv 32768 on 2x2x2) incl. a bc sweep, cfg4 (SUMMA 65536 on 2x2x2),
# 3D CA-CQR2 on the cube with the device validators, grid parity at 8 ranks.  Every step under its own timeout.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
B=capital_b200/host/bin
timeout 300 $TR --master-port 29601 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
tail -c 2500 gpurun_out/bench_n8.json; tail -3 gpurun_out/bench_n8.err
timeout 200 $TR --master-port 29602 tools/grid_check.py 2048 > gpurun_out/grid_check_8.log 2>&1; tail -12 gpurun_out/grid_check_8.log
export CAPITAL_B200_VALIDATE=1
timeout 200 $TR --master-port 29603 --no-python $B/cholinv 32768 1 1 1 -2 0 0 3 > gpurun_out/cfg3_ci1.log 2>&1; grep -E "total time|residual|rror" gpurun_out/cfg3_ci1.log | tail -4
timeout 200 $TR --master-port 29604 --no-python $B/cholinv 32768 1 0 1 -2 0 0 3 > gpurun_out/cfg3_ci0.log 2>&1; grep -E "total time|residual|rror" gpurun_out/cfg3_ci0.log | tail -4
timeout 300 $TR --master-port 29605 --no-python $B/cacqr 2 1048576 4096 2 2 1 1 -2 -2 0 0 2 > gpurun_out/cacqr3d.log 2>&1; grep -E "TFLOP|orthogonality|rror" gpurun_out/cacqr3d.log | tail -4
unset CAPITAL_B200_VALIDATE
timeout 300 $TR --master-port 29606 --no-python $B/summa_gemm 65536 65536 65536 1 0 0 2 > gpurun_out/cfg4.log 2>&1; grep -E "TFLOP|rror" gpurun_out/cfg4.log | tail -3
timeout 300 $TR --master-port 29607 tools/sweep.py cholinv --N 32768 --bc -1 -2 -3 -4 --ci 0 --iters 2 --out gpurun_out/sweep_cfg3.json > gpurun_out/sweep_cfg3.log 2>&1; tail -8 gpurun_out/sweep_cfg3.log
