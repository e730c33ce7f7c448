#!/bin/bash
# last short 1-GPU session of the round: the full GPU suite, then the bench line if time is left
mkdir -p gpurun_out
S=$(date +%s)
timeout 125 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$? after $(( $(date +%s) - S )) s"; tail -4 gpurun_out/pytest_gpu_final.log
timeout 100 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1_short.json 2> gpurun_out/bench_n1_short.err; echo "bench rc=$? after $(( $(date +%s) - S )) s"; tail -c 2500 gpurun_out/bench_n1_short.json
