#!/bin/bash
# A/B of the front end's buffer-set count (GFORB_FE_NB), alternating so that box-to-box and run-to-run drift cancels
run() { python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --no-weak 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'value',round(d['value']),'e2e',round(d['e2e']['value']), d['roofline']['pipeline_ms'])"; }
for i in 1 2 3; do
GFORB_FE_NB=2 run nb2
GFORB_FE_NB=3 run nb3
GFORB_FE_NB=4 run nb4
done
