#!/bin/bash
# pipelined SUMMA A/B at P ranks: cholinv N=32768 (reference schedule) with the serial and the pipelined panel loops
P=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
run() { echo "== $*"; env "$@" timeout 200 $TR --master-port 29642 tools/sweep.py cholinv --N ${N:-32768} --bc -2 --ci 0 --iters 2 --out gpurun_out/sweep_tmp.json 2>&1 | grep -E "^cholinv" ; }
run CAPITAL_B200_SUMMA_PIPE=0
run CAPITAL_B200_SUMMA_PIPE=1 CAPITAL_B200_SUMMA_RESERVE=0
run CAPITAL_B200_SUMMA_PIPE=1 CAPITAL_B200_SUMMA_RESERVE=12
run CAPITAL_B200_SUMMA_PIPE=1 CAPITAL_B200_SUMMA_RESERVE=24
run CAPITAL_B200_SUMMA_PIPE=1 CAPITAL_B200_SUMMA_RESERVE=12 CAPITAL_B200_SUMMA_HALVES=0
run CAPITAL_B200_SUMMA_PIPE=1 CAPITAL_B200_SUMMA_RESERVE=12 NCCL_MAX_CTAS=4
