#!/bin/bash
# quick GPU iteration: parity tests + kernel timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python tools/kernel_probe.py "$@" > gpurun_out/probe.log 2>&1
tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/probe.log
