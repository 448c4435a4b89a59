#!/bin/bash
# alternating A/B runs of bench.py under environment knobs (box-to-box and run-to-run drift cancels): usage gpu_r2_ab.sh "A=1 B=2" "C=3" ...
run() { env $2 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'value',round(d['value']),'e2e',round(d['e2e']['value']), d['roofline']['pipeline_ms'])"; }
for i in 1 2 3; do
  run base "X=0"
  n=0
  for v in "$@"; do n=$((n+1)); run "v$n[$v]" "$v"; done
done
