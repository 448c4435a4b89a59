#!/bin/bash
# A/B of the blur fork inside gforb_extract (GFORB_BLUR_FORK=0/1): parity tests, then the bench without the CPU legs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extractor.py tests/test_gpu_pipeline.py -m gpu -x -q 2>&1 | tail -5
for f in 0 1; do
  GFORB_BLUR_FORK=$f timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak > gpurun_out/bench_fork$f.json 2> gpurun_out/bench_fork$f.err
  echo "fork=$f rc=$?"
  python -c "
import json
d=json.loads(open('gpurun_out/bench_fork$f.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
print(d['roofline']['pipeline_ms'], d['roofline']['pipeline_ms_serial'])"
done
