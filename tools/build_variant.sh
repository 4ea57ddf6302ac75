#!/bin/bash
# Builds build/var_<name>/libfastbn_b200.so with extra -D flags (tuning experiments; selected at run time with
# FASTBN_B200_LIB, see fastbn_b200/capi.py):   tools/build_variant.sh <name> -DFBN_FLAT_OCC=32 -DFBN_FLAT_BATCH=2 ...
# Every CUDA file and the planner are rebuilt (plan.h constants such as FBN_W reach all of them); host objects are reused.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=build/var_$name; mkdir -p $out
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fopenmp -Xptxas -v"
for f in kernels level_kernels flat_kernel resident_kernel noalloc capi comm; do
  $NV "$@" -c fastbn_b200/csrc/$f.cu -o $out/$f.o 2> $out/$f.ptxas.log &
done
g++ -O2 -std=c++17 -fPIC "$@" -c fastbn_b200/csrc/plan.cpp -o $out/plan.o &
g++ -O2 -std=c++17 -fPIC "$@" -c fastbn_b200/csrc/resident.cpp -o $out/resident.o &
wait
objs=$(ls build/csrc/host_*.o)
$NV -shared -o $out/libfastbn_b200.so $objs $out/plan.o $out/resident.o $out/kernels.o $out/level_kernels.o $out/flat_kernel.o $out/resident_kernel.o $out/noalloc.o $out/capi.o $out/comm.o -Xcompiler -fopenmp -lgomp 2>/dev/null
grep -h -B2 "Used" $out/flat_kernel.ptxas.log | grep -E "spill|Used" | tr '\n' ' '; echo
echo built $out/libfastbn_b200.so
