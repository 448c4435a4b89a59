#!/bin/bash
# quick pass: extractor / pipeline parity tests, then two bench runs (serial stage times + pipelined frames/s)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extractor.py tests/test_gpu_pipeline.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2_q2_pytest.log
run() { env $2 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'arm':'$1','value':round(d['value']),'e2e':round(d['e2e']['value']),'stage_ms_serial':d['roofline']['stage_ms_serial'],'pipeline_ms_serial':d['roofline']['pipeline_ms_serial'],'pipeline_ms':d['roofline']['pipeline_ms']}))"; }
for i in 1 2; do run new "X=0"; for v in "$@"; do run "[$v]" "$v"; done; done | tee gpurun_out/r2_q2_ab.jsonl
