#!/bin/bash
# round-2 extractor kernel generations: parity tests, alternating bench A/B (serial stage times + pipelined frames/s), ncu --set full
# of the kernels that changed (one launch each, after the plain command ran clean)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extractor.py tests/test_gpu_pipeline.py tests/test_gpu_shim.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2_kp_pytest.log
run() { env $2 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'arm':'$1','value':round(d['value']),'e2e':round(d['e2e']['value']),'stage_ms_serial':d['roofline']['stage_ms_serial'],'pipeline_ms':d['roofline']['pipeline_ms']}))"; }
for i in 1 2 3; do
  run old "GFORB_KP=v1 GFORB_BLUR=v2 GFORB_FN_TALL=0"
  run new "X=0"
  if [ $i = 1 ]; then run new_short_strips "GFORB_FN_TALL=0"; run new_blur_v2 "GFORB_BLUR=v2"; run new_describe_smem "GFORB_DESCRIBE=smem"; fi
done | tee gpurun_out/r2_kp_ab.jsonl
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra --no-weak --seq-frames 128 --batch 128"
$CMD > gpurun_out/plain_kp.log 2>&1; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_describe|k_orient2|k_blur3|k_fast_nms" -s 12 -c 4 -o gpurun_out/r2_kp_new -f $CMD > gpurun_out/ncu_kp2.log 2>&1; tail -2 gpurun_out/ncu_kp2.log
GFORB_DESCRIBE=smem GFORB_BLUR=v2 GFORB_FN_TALL=0 timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_describe2|k_blur2|k_fast_nms" -s 9 -c 3 -o gpurun_out/r2_kp_alt -f $CMD > gpurun_out/ncu_kp1.log 2>&1; tail -2 gpurun_out/ncu_kp1.log
ls -la gpurun_out | tail -8
