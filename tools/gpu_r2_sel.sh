#!/bin/bash
mkdir -p gpurun_out
for b in 0 1; do echo "BAND=$b"; GFORB_SELECT_BAND=$b python bench.py --steps 6 --warmup 3 --no-cpu --no-extra 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['pipeline_ms'], d['roofline']['pipeline_ms_serial'])"; done
