python -m pytest tests/test_gpu_shapes.py -m gpu -x -q 2>&1 | tail -2
for a in "pigs 58944 0 2 2" "munin_small 3040 0 2 0" "munin 1216 0 2 0"; do
  timeout 600 python tools/prof_run.py $a 2>&1 | tail -1
done
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r1d_launches_munin.csv python tools/prof_run.py munin 608 0 1 0 > gpurun_out/ncu_l.log 2>&1
