#!/bin/bash
# multi-GPU session for cap_potrf_1d.  usage: gpu_potrf1d_multi.sh P TIMEOUT "args of potrf1d_check.py" ...
P=$1; T=$2; shift 2
mkdir -p gpurun_out
for cfg in "$@"; do
  timeout $T python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29571 tools/potrf1d_check.py $cfg --json gpurun_out/potrf1d_n$P.jsonl 2>&1 | grep -E "potrf_1d|rank|Error|error|Traceback|NCCL WARN" | head -40
  echo "rc=${PIPESTATUS[0]} for: $cfg"
done
