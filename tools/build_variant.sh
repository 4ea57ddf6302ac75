#!/bin/bash
# Builds build/var_<name>/libfastbn_b200.so with extra -D flags for the kernel files (tuning experiments; selected at run
# time with FASTBN_B200_LIB, see fastbn_b200/capi.py):   tools/build_variant.sh <name> -DFBN_PULL_BATCH=8 ...
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=build/var_$name; mkdir -p $out
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fopenmp"
for f in kernels level_kernels flat_kernel; do
  $NV "$@" -c fastbn_b200/csrc/$f.cu -o $out/$f.o &
done
wait
objs=$(ls build/csrc/*.o | grep -v -E "/(kernels|level_kernels|flat_kernel)\.o")
$NV -shared -o $out/libfastbn_b200.so $objs $out/kernels.o $out/level_kernels.o $out/flat_kernel.o -Xcompiler -fopenmp -lgomp
echo built $out/libfastbn_b200.so
