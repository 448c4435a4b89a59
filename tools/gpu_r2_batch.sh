#!/bin/bash
# frames per front-end batch: 250 (default) vs 500 vs 1000 at N = 1
mkdir -p gpurun_out
for B in 250 500 1000; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak --batch $B > gpurun_out/bench_b$B.json 2> gpurun_out/bench_b$B.err
  echo "batch=$B rc=$?"; tail -2 gpurun_out/bench_b$B.err
  python -c "
import json
d=json.loads(open('gpurun_out/bench_b$B.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
print(d['roofline']['pipeline_ms'], d['roofline']['pipeline_ms_serial'])"
done
