#!/bin/bash
# experiment: four test cases per lane (FBN_W=4: 1 KB rows, 128-case tiles) in the fused / level kernels
cd "$(dirname "$0")/.."
run() { echo "== $*"; timeout 600 python tools/prof_run.py "$@" | tail -1; }
unset FASTBN_B200_LIB
run hepar2 200000 8 2 1
run pigs 60000 16 2 1
run munin 2048 0 2 2
export FASTBN_B200_LIB=$PWD/build/var_w4/libfastbn_b200.so
FBN_CB=2 run hepar2 200000 8 2 1
FBN_CB=2 run hepar2 200000 4 2 1
FBN_CB=2 run pigs 60000 16 2 1
run munin 2048 0 2 2
export FASTBN_B200_LIB=$PWD/build/var_w4o8/libfastbn_b200.so
run hepar2 200000 8 2 1
run hepar2 200000 4 2 1
FBN_CB=2 run hepar2 200000 8 2 1
run pigs 60000 8 2 1
