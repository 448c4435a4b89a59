#!/bin/bash
# 8-GPU pass: bench.py under torchrun (incl. the all-ranks-copying H2D ceiling of the e2e block) and config 5 sharded
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
tail -3 gpurun_out/bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e'])
print('weak',d.get('weak_scaling')); print('C5',d.get('extra',{}).get('C5'))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/run_c5.py > gpurun_out/c5_n$N.json 2> gpurun_out/c5_n$N.err; echo "c5 N=$N rc=$?"; tail -1 gpurun_out/c5_n$N.json | cut -c1-600
