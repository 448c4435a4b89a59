#!/bin/bash
# 8-GPU end-to-end diagnosis: bench without the CPU / extra / weak legs, e2e stage times per batch
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-extra --no-weak > gpurun_out/bench_n${N}b.json 2> gpurun_out/bench_n${N}b.err; echo "bench N=$N rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n${N}b.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e'])
PY
