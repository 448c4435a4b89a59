#!/bin/bash
# final round-2 evidence pass, third edition (k_resize2, k_blur3 with two rows in flight, one-round row mapping) on one B200: whole GPU test suite, smoke, bench (our arm + reference arm), ncu launch list and one
# --set full capture of every kernel of a 128-frame step (each only after its own command ran clean without ncu)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2_final3_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r2_final3_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_final3_bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_final3_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'cpu',d.get('cpu_baseline',{}).get('value'),'frac',d['roofline']['frac'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_final3_bench_reference_arm.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/r2_final3_bench_reference_arm.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra --no-weak --seq-frames 128 --batch 128"
$CMD > gpurun_out/plain_prof.log 2>&1; echo "plain rc=$?"
LPS=$(python -c "import json;d=json.loads(open('gpurun_out/plain_prof.log').read().strip().splitlines()[-1]);print(d['gpu_launches']//d['steps'])")
echo LPS=$LPS
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_final3_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
tail -2 gpurun_out/ncu_launch.log
timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:^k_" -s $((3 * LPS)) -c $LPS -o gpurun_out/r2_final3_step_full -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/ | tail -12
