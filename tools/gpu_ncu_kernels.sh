#!/bin/bash
# ncu --set full capture of kernels matching a regex inside one front-end step.
# usage: gpu_ncu_kernels.sh <outname> <regex> [count] [batch]
set -x
mkdir -p gpurun_out
B=${4:-128}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --batch $B --distinct 16"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:$2" -s ${5:-0} -c ${3:-4} -o gpurun_out/$1 -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
