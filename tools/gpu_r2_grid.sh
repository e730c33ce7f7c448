#!/bin/bash
P=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29641 tools/grid_check.py ${2:-1024} 2>&1 | grep -E "rel err|GRID CHECK|validate|rror" | tail -40
timeout 200 $TR --master-port 29643 --no-python capital_b200/host/bin/cholinv 4096 1 1 1 -2 1 3 1 2>&1 | grep -E "total time|residual|rror" | tail -3
