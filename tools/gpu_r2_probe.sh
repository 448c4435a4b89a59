#!/bin/bash
# round-2 probe: tcgen05 descriptor / throughput probe, integer-pipe peaks, then the GPU test-suite as a baseline
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 120 tools/ubench/tc_probe.bin > gpurun_out/tc_probe.jsonl 2>&1; echo "tc_probe rc=$?"
cat gpurun_out/tc_probe.jsonl
timeout 120 tools/ubench/pipes.bin > gpurun_out/int_pipe.jsonl 2>&1; echo "pipes rc=$?"
cat gpurun_out/int_pipe.jsonl
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.log
