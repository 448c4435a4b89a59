#!/bin/bash
# Full GPU evidence pass: gpu tests, smoke, bench, ncu launch list, ncu --set full of one step.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.log
bash tools/gpu_round.sh
LPS=$(python -c "import json;d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1]);print(d['gpu_launches']//d['steps'])")
echo LPS=$LPS
timeout 1200 bash tools/gpu_ncu_step.sh ${1:-step_full} 128 $LPS
