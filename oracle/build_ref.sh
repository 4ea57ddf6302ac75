#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Compiles the UNMODIFIED reference from the sources where they lie under
# /root/reference (no copies) plus oracle/ref_harness.cpp into oracle/_ref/libfastbn_ref.so.
# Does not use the reference's own CMake build. Workarounds are compiler flags only:
#   -include cstdint -include climits : g++ 13 needs <cstdint>/<climits> for INT32_MIN / INT_MAX (include/common.h:30,36)
# oracle/_ref/ is git-ignored but travels to the GPU box with the snapshot.
set -euo pipefail
REF=${FASTBN_REFERENCE:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
  echo "build_ref.sh: $REF not present; keeping prebuilt $OUT if any" >&2
  exit 0
fi
mkdir -p "$OUT/obj"
CXX=${CXX_REF:-/usr/bin/g++}
FLAGS="-O3 -std=c++14 -fopenmp -fPIC -w -include cstdint -include climits -I$REF/include -I$REF/lib/tinyxml2 -I$REF/lib/ARFF/src -I$REF/lib/stats/include -I$REF/lib/gcem/include"
SRCS=$(ls "$REF"/src/*.cpp | grep -v '/main.cpp$')
OBJS=""
pids=()
for s in $SRCS "$REF/lib/tinyxml2/tinyxml2.cpp" "$HERE/ref_harness.cpp"; do
  o="$OUT/obj/$(basename "${s%.cpp}").o"
  OBJS="$OBJS $o"
  if [ ! -f "$o" ] || [ "$s" -nt "$o" ]; then
    $CXX $FLAGS -c "$s" -o "$o" &
    pids+=($!)
    if [ ${#pids[@]} -ge 8 ]; then wait "${pids[0]}"; pids=("${pids[@]:1}"); fi
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
$CXX -shared -fopenmp -o "$OUT/libfastbn_ref.so" $OBJS
echo "built $OUT/libfastbn_ref.so"
