#!/bin/bash
# ncu --set full capture of every gforb kernel of ONE front-end step (after 3 warm-up steps).
# usage: gpu_ncu_step.sh <outname> [batch] [launches_per_step]
set -x
mkdir -p gpurun_out
B=${2:-128}
LPS=${3:-20}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --batch $B --distinct 16"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:^k_" -s $((3 * LPS)) -c $LPS -o gpurun_out/$1 -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/
