#!/bin/bash
# 1-GPU session: look-ahead Cholesky parity (one rank, both pipeline modes) + timing against the recursive schedule
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_solve.py -q -m gpu -k "lookahead" > gpurun_out/pytest_lookahead.log 2>&1; echo "pytest lookahead rc=$?"; tail -5 gpurun_out/pytest_lookahead.log
timeout 200 python tools/potrf1d_check.py 16384 1024 2048 --time 3 --reserve 8 16 --pipeline 0 1 --json gpurun_out/potrf1d_n1.jsonl 2>&1 | grep -E "potrf_1d|rank|Error|error"
timeout 100 python tools/potrf1d_check.py 8192 1024 --time 3 --reserve 8 --pipeline 0 1 --json gpurun_out/potrf1d_n1.jsonl 2>&1 | grep -E "potrf_1d|rank|Error|error"
