#!/bin/bash
# 1-GPU session: look-ahead Cholesky parity (one rank) + timing against the recursive schedule
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_solve.py -q -m gpu -k "lookahead" > gpurun_out/pytest_lookahead.log 2>&1; echo "pytest lookahead rc=$?"; tail -25 gpurun_out/pytest_lookahead.log
for cfg in "8192 1024" "16384 1024" "16384 512" "16384 2048"; do
  timeout 120 python tools/potrf1d_check.py $cfg --time 3 --reserve 0 8 16 24 --json gpurun_out/potrf1d_n1.jsonl 2>&1 | grep -E "potrf_1d|rank|Error|error"
done
