#!/bin/bash
# durations of the matcher kernels at batch 250 (ncu launch list; lane-parallel and speculative INIT resolver)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-extra --no-weak --seq-frames 250 --batch 250"
for v in lanes spec; do
  GFORB_RESOLVE=$v timeout 600 ncu --metrics gpu__time_duration.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -k "regex:k_resolve|k_candidates|k_grid_build|k_fe_prep|k_build_info" -c 20 --csv --log-file gpurun_out/match_$v.csv $CMD > /dev/null 2>&1
  echo "== $v"; python - <<PY
import csv
rows=[r for r in csv.reader(l for l in open('gpurun_out/match_$v.csv') if not l.startswith('=='))]
h=rows[0]; ki=h.index('Kernel Name'); mi=h.index('Metric Name'); vi=h.index('Metric Value'); ii=h.index('ID')
d={}
for r in rows[1:]:
    d.setdefault((r[ii], r[ki][:40]),{})[r[mi]]=r[vi]
for (i,k),m in list(d.items())[:10]: print(i,k,m)
PY
done
