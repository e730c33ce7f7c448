#!/bin/bash
# ncu --set full of the timed launches of the given probe cases (plain run first, as the recipe requires)
mkdir -p gpurun_out
NAME=$1; shift
P="python tools/kernel_probe.py $* --reps 1 --profile"
$P > gpurun_out/plain_$NAME.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/prof_$NAME $P > gpurun_out/ncu_$NAME.log 2>&1
cat gpurun_out/plain_$NAME.log; tail -3 gpurun_out/ncu_$NAME.log
