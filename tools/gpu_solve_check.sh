#!/bin/bash
# short 1-GPU session: parity of cap_dtrsm / blocked cap_dpotrf / batched-leaf cap_dtrtri + the cfg5 probe
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_solve.py -q -m gpu > gpurun_out/pytest_solve.log 2>&1; echo "pytest solve rc=$?"; tail -25 gpurun_out/pytest_solve.log
timeout 120 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "potrf or trtri or cholinv" > gpurun_out/pytest_factor.log 2>&1; echo "pytest factor rc=$?"; tail -5 gpurun_out/pytest_factor.log
timeout 200 python tools/factor_probe.py --json gpurun_out/factor_probe.json > gpurun_out/factor_probe.log 2>&1; echo "probe rc=$?"; cat gpurun_out/factor_probe.log
