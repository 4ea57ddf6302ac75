#!/bin/bash
# GPU run of the resident-tile kernel: smoke (with a timeout: a wrong mbarrier phase would hang), warps-per-CTA / grain sweep
# against the flat kernel, the GPU tests
cd "$(dirname "$0")/.."
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2n_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2n_smoke.log
for w in ${WARPS:-8 16 32}; do for g in ${GRAINS:-60 120}; do echo "== resident warps $w grain $g"; FBN_RES_GRAIN=$g FBN_RES_WARPS=$w timeout 300 python tools/prof_run.py alarm 1000000 0 3 | tail -1; done; done
true
echo "== flat"; FBN_NO_RESIDENT=1 timeout 300 python tools/prof_run.py alarm 1000000 0 3 | tail -1
if [ -z "$NOTESTS" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2n_gpu_tests.log; fi
