#!/bin/bash
# ncu launch lists (device time + DRAM bytes per launch) of one device-resident pass of every workload: the source of
# profiles/traffic.json (numbers printed under ncu are never bench values). Each capture follows a plain run of the same command.
cd "$(dirname "$0")/.."
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
tag=${1:-r2}
for spec in "alarm 1000000" "hepar2 200000" "hepar2_50 200000" "pigs 60000" "munin 2048"; do
  set -- $spec
  python tools/prof_run.py $1 $2 0 2 > gpurun_out/${tag}_plain_$1.log 2>&1 &&
  ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_$1.csv python tools/prof_run.py $1 $2 0 1 > gpurun_out/${tag}_ncu_$1.log 2>&1
  echo "$1 rc=$?"
done
python tools/update_traffic.py $tag gpurun_out && cp profiles/traffic.json gpurun_out/${tag}_traffic.json
