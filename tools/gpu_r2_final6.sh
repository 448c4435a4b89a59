#!/bin/bash
# last evidence pass of round 2 (HEAD of round 2: default batches of <= 296 frames, k_cell_emit without the division; kernels otherwise as profiled in r2_final3_*): whole GPU test suite, smoke,
# bench (our arm + reference arm)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2_final6_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r2_final6_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_final6_bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_final6_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'cpu',d.get('cpu_baseline',{}).get('value'),'frac',d['roofline']['frac'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_final6_bench_reference_arm.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/r2_final6_bench_reference_arm.json
