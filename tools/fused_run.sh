#!/bin/bash
# one device-resident pass of the fused / level workloads (plain runs, CUDA-event timing): quick check after a kernel change
cd "$(dirname "$0")/.."
run() { echo "== $*"; timeout 600 python tools/prof_run.py "$@" | tail -1; }
run hepar2 200000 0 2
run hepar2_50 200000 0 2
run pigs 60000 0 2
run munin 2048 0 2
if [ -z "$NOTESTS" ]; then timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
