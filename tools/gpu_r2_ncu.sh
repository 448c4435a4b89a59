#!/bin/bash
# ncu --set full captures (with source) of the tensor-core all-pairs kernel and the selection kernel, one launch each
mkdir -p gpurun_out
C1="python tools/bench_allpairs.py --only tc --iters 1"
$C1 > gpurun_out/plain_tc.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_allpairs_tc -s 1 -c 1 -o gpurun_out/r2_allpairs_tc -f $C1 > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
C2="python tools/bench_select.py 256 1"
$C2 > gpurun_out/plain_sel.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_select_ahead -s 2 -c 1 -o gpurun_out/r2_select -f $C2 > gpurun_out/ncu_sel.log 2>&1
tail -2 gpurun_out/ncu_sel.log; cat gpurun_out/plain_sel.log | tail -1
ls -la gpurun_out/*.ncu-rep
