#!/bin/bash
mkdir -p gpurun_out
python tools/kernel_probe.py update --N 8192 --K 256 512 1024 2048 4096 --reps 3 2>&1 | grep -E "update|rror"
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
timeout 200 python tools/potrf1d_check.py 16384 1024 2048 --time 3 --pipeline -1 --reserve 16 --grid 64 2>&1 | grep -E "^potrf_1d|FAIL|rror"
timeout 200 python tools/potrf1d_check.py 32768 2048 1024 --time 2 --pipeline -1 --reserve 16 --grid 64 2>&1 | grep -E "^potrf_1d|FAIL|rror"
