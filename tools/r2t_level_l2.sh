#!/bin/bash
# experiment: HEPAR2 / PIGS shapes in level mode with so few tiles in flight that their messages stay in the 126 MB L2
cd "$(dirname "$0")/.."
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
for r in 4 8 16; do
  echo "== hepar2 level resident $r"; FBN_LEVEL_RESIDENT=$r python tools/prof_run.py hepar2 8192 0 2 2 | tail -1
done
FBN_LEVEL_RESIDENT=8 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r2t_launches_hepar2_level8.csv python tools/prof_run.py hepar2 1024 0 1 2 > gpurun_out/r2t_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/r2t_launches_hepar2_level8.csv 1024 4
echo "== pigs level resident 4"; FBN_LEVEL_RESIDENT=4 python tools/prof_run.py pigs 4096 0 2 2 | tail -1
