#!/bin/bash
mkdir -p gpurun_out
python tools/kernel_probe.py update --N 8192 --K 512 1024 2048 --reps 3 2>&1 | grep -E "update|rror"
python tools/kernel_probe.py gemm_tn apply gram --reps 2 2>&1 | grep -E "TFLOP|rror"
