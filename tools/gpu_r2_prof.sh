#!/bin/bash
# round-2 ncu evidence of the front-end step: launch list (durations only) and one --set full capture of a 128-frame batch
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra --no-weak --seq-frames 128 --batch 128"
$CMD > gpurun_out/plain_prof.log 2>&1; echo "plain rc=$?"
LPS=$(python -c "import json;d=json.loads(open('gpurun_out/plain_prof.log').read().strip().splitlines()[-1]);print(d['gpu_launches']//d['steps'])")
echo LPS=$LPS
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
tail -2 gpurun_out/ncu_launch.log
timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:^k_" -s $((3 * LPS)) -c $LPS -o gpurun_out/r2_step_full -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/ | tail -8
