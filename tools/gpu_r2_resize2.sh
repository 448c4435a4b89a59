#!/bin/bash
# k_resize2 (GFORB_RESIZE=words) against the byte-gather k_resize: parity tests, alternating bench A/B, ncu --set full of both
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extractor.py -m gpu -x -q -k "resize or extract_matches" 2>&1 | tail -4 | tee gpurun_out/r2_rs2_pytest.log
run() { env $2 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'arm':'$1','value':round(d['value']),'e2e':round(d['e2e']['value']),'stage_ms_serial':d['roofline']['stage_ms_serial'],'pipeline_ms':d['roofline']['pipeline_ms']}))"; }
for i in 1 2 3; do run direct "GFORB_RESIZE=direct"; run words "GFORB_RESIZE=words"; run words8 "GFORB_RESIZE=words8"; done | tee gpurun_out/r2_rs2_ab.jsonl
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-extra --no-weak --seq-frames 128 --batch 128"
GFORB_RESIZE=words $CMD > gpurun_out/plain_rs2.log 2>&1; echo "plain rc=$?"
GFORB_RESIZE=words timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_resize2" -s 21 -c 7 -o gpurun_out/r2_rs2_words4 -f $CMD > gpurun_out/ncu_rs2.log 2>&1; tail -2 gpurun_out/ncu_rs2.log
GFORB_RESIZE=words8 timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_resize2" -s 21 -c 7 -o gpurun_out/r2_rs2_words8 -f $CMD > gpurun_out/ncu_rs2b.log 2>&1; tail -2 gpurun_out/ncu_rs2b.log
