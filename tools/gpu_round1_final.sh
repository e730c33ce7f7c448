#!/bin/bash
# final 1-GPU session of the round: full GPU test suite, bench line, numbers for the BASELINE.md table,
# and the ncu launch list of the bench command (after it exited 0 without ncu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_final.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_n1.json
timeout 300 python tools/kernel_probe.py cholinv --n 16384 --reps 2 2>&1 | grep -E "cholinv n"
timeout 300 python tools/kernel_probe.py potrf trtri gemm_nn apply gram --m 4194304 --reps 2 2>&1 | grep -E "potrf|trtri|gemm|apply|gram"
capital_b200/host/bin/cacqr 2 65536 256 1 1 1 1 0 0 0 0 3 | tail -1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29631 tools/sweep.py cholinv --N 4096 --bc -1 -2 --ci 0 --iters 1 2>&1 | tail -4
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_r01.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"; tail -c 300 gpurun_out/ncu_bench.log
