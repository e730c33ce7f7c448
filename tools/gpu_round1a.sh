#!/bin/bash
# first GPU pass of the session: parity tests, peak calibration, kernel timings, ncu evidence
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
./tools/dmma_peak > gpurun_out/dmma_peak.log 2>&1
python tools/kernel_probe.py gemm_nn gemm_tn gemm_nt apply apply_inplace gram cholinv cqr2 potrf trtri --reps 3 > gpurun_out/probe.log 2>&1
P="python tools/kernel_probe.py gemm_nn apply gram --reps 1 --m 524288 --N 4096 --profile"
$P > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/prof_r01_gemm $P > gpurun_out/ncu_full.log 2>&1
P2="python tools/kernel_probe.py cqr2 --reps 1 --m 524288 --profile"
$P2 > gpurun_out/plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_cqr2.csv $P2 > gpurun_out/ncu_list.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/dmma_peak.log gpurun_out/probe.log
