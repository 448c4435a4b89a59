#!/bin/bash
# ncu --set full capture of selected kernels (one GPU).  usage: gpu_ncu_full.sh <regex> <count> <outname>
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --batch 64 --distinct 16"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:$1" -s ${4:-0} -c $2 -o gpurun_out/$3 -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/
