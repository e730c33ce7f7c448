#!/bin/bash
# First multi-GPU session of round 2: where does the time of cap_potrf_1d go at 8 GPUs,
# and does keeping NCCL's spinning kernels off the reserved SMs un-stick the pipelined order?
#   gpurun --gpus 8 --timeout 400 -- 'bash tools/gpu_round2_first.sh 8'
# Every line of tools/potrf1d_check.py is checked against the single-GPU factor; timelines go to gpurun_out/.
P=${1:-8}; N=${2:-32768}; NB=${3:-1024}
mkdir -p gpurun_out
run() {  # run <tag> <extra args...>; environment of the caller is inherited (NCCL_MAX_CTAS)
  tag=$1; shift
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29571 \
    tools/potrf1d_check.py $N $NB --time 3 --reserve 16 --json gpurun_out/potrf1d_r02_n$P.jsonl "$@" \
    > gpurun_out/potrf1d_$tag.out 2> gpurun_out/potrf1d_$tag.err
  echo "== $tag rc=$?"; grep -E "^potrf_1d|FAIL" gpurun_out/potrf1d_$tag.out
}
run base --pipeline 0 1
NCCL_MAX_CTAS=4 run ctas4 --pipeline 0 1
NCCL_MAX_CTAS=2 run ctas2 --pipeline 0 1
run timeline_p0 --pipeline 0 --timeline      # per-panel phase timeline of every rank (warm call) on stderr
run timeline_p1 --pipeline 1 --timeline
grep -c "timeline" gpurun_out/potrf1d_timeline_p0.err gpurun_out/potrf1d_timeline_p1.err
