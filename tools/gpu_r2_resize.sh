#!/bin/bash
# A/B of the pyramid resize: gathers from L1 (k_resize) vs source rows staged by cp.async.bulk (k_resize_tma)
mkdir -p gpurun_out
for v in direct tma; do
  echo "GFORB_RESIZE=$v"
  GFORB_RESIZE=$v timeout 600 python -m pytest tests/test_gpu_extractor.py -m gpu -x -q 2>&1 | tail -2
  GFORB_RESIZE=$v python bench.py --steps 6 --warmup 3 --no-cpu --no-extra 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['stage_ms_serial'])"
done
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --seq-frames 250"
for v in direct tma; do
  GFORB_RESIZE=$v $CMD > gpurun_out/plain_resize_$v.log 2>&1 && \
  GFORB_RESIZE=$v ncu --set full --clock-control none -k regex:k_resize -s 14 -c 2 -o gpurun_out/r2_resize_$v -f $CMD > gpurun_out/ncu_resize_$v.log 2>&1
  tail -1 gpurun_out/ncu_resize_$v.log
done
