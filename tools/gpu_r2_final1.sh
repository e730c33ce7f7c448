#!/bin/bash
mkdir -p gpurun_out
echo "== new test alone (fresh process)"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/r02_pytest_noalloc_alone.txt
echo "== whole GPU suite"; timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | tee gpurun_out/r02_pytest_gpu.txt
timeout 500 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_n1_final.json 2> gpurun_out/r02_bench_n1_final.err; echo bench_rc=$?; tail -c 400 gpurun_out/r02_bench_n1_final.err
