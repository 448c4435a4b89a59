#!/bin/bash
# parity tests of the front-end kernels + a bench without the CPU legs
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_extractor.py tests/test_gpu_matcher.py tests/test_gpu_pipeline.py tests/test_gpu_shim.py -m gpu -x -q 2>&1 | tail -8
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
echo "bench rc=$?"; tail -3 gpurun_out/bench_quick.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
print({k:v['ms'] for k,v in d['roofline']['per_stage_serial'].items()})
print(d['roofline']['pipeline_ms'], d['roofline']['pipeline_ms_serial'])"
