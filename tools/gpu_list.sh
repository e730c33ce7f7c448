#!/bin/bash
# ncu launch list (device time per launch) of the timed launches of the given probe cases
mkdir -p gpurun_out
NAME=$1; shift
P="python tools/kernel_probe.py $* --reps 1 --profile"
$P > gpurun_out/plain_$NAME.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_$NAME.csv $P > gpurun_out/nculist_$NAME.log 2>&1
cat gpurun_out/plain_$NAME.log
