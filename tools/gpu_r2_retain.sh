#!/bin/bash
# A/B of retainBest's nth_element: one lane (GFORB_RETAIN=seq) vs the warp-parallel replay (default); parity tests first
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extractor.py tests/test_gpu_pipeline.py -m gpu -x -q 2>&1 | tail -8
for f in seq warp; do
  GFORB_RETAIN=$f timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak > gpurun_out/bench_retain_$f.json 2> gpurun_out/bench_retain_$f.err
  echo "retain=$f rc=$?"
  python -c "
import json
d=json.loads(open('gpurun_out/bench_retain_$f.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
print({k:v['ms'] for k,v in d['roofline']['per_stage_serial'].items()})
print(d['roofline']['pipeline_ms'], d['roofline']['pipeline_ms_serial'])"
done
