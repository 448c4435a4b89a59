#!/bin/bash
# selection kernel alone: 250 C4 problems, phase cycle counters of problem 0 for the thread-count / residency variants
cd /root/repo
for cfg in "128 2" "128 3" "256 1" "64 5"; do
  set -- $cfg
  echo "== T=$1 MINB=$2"
  GFORB_SELECT_PROF=1 GFORB_SELECT_THREADS=$1 GFORB_SELECT_MINB=$2 timeout 300 python tools/bench_select.py 250 3 2>&1 | grep -v "^$" | tail -3
done
echo "== batch 32 (latency of one wave, T=128)"; GFORB_SELECT_PROF=1 GFORB_SELECT_THREADS=128 timeout 300 python tools/bench_select.py 32 3 2>&1 | tail -2
