#!/bin/bash
# Round-2 tuning sweep (plain runs, CUDA-event timing inside tools/prof_run.py): flat-kernel variants on ALARM, team sizes on
# the HEPAR2 / PIGS shapes, resident tiles on the MUNIN shape.
cd "$(dirname "$0")/.."
for v in default ${VARIANTS}; do
  if [ "$v" != default ]; then export FASTBN_B200_LIB=$PWD/build/var_$v/libfastbn_b200.so; else unset FASTBN_B200_LIB; fi
  echo "== flat variant $v"; python tools/prof_run.py alarm 1000000 0 3 0 | tail -1
done
unset FASTBN_B200_LIB
for t in ${HEPAR_TEAMS:-2 4 8}; do echo "== hepar2 team $t"; python tools/prof_run.py hepar2 200000 $t 2 1 | tail -1; done
for t in ${PIGS_TEAMS:-4 8 16}; do echo "== pigs team $t"; python tools/prof_run.py pigs 60000 $t 2 1 | tail -1; done
for r in ${MUNIN_RES:-8 32 128}; do echo "== munin resident $r"; FBN_LEVEL_RESIDENT=$r python tools/prof_run.py munin 2048 0 2 2 | tail -1; done
