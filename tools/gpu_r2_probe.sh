#!/bin/bash
# round-2 probe: tcgen05 probe, TC all-pairs tests + A/B bench, then the whole GPU test-suite
mkdir -p gpurun_out
timeout 120 tools/ubench/tc_probe.bin > gpurun_out/tc_probe.jsonl 2>&1; echo "tc_probe rc=$?"
cat gpurun_out/tc_probe.jsonl
timeout 600 python -m pytest tests/test_gpu_matcher.py -m gpu -x -q -k allpairs 2>&1 | tail -15
timeout 300 python tools/bench_allpairs.py > gpurun_out/allpairs_ab.json 2> gpurun_out/allpairs_ab.err; echo "ab rc=$?"; cat gpurun_out/allpairs_ab.json; tail -3 gpurun_out/allpairs_ab.err
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
