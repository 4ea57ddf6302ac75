for v in default occ4; do
  if [ "$v" != default ]; then export FASTBN_B200_LIB=$PWD/build/var_$v/libfastbn_b200.so; else unset FASTBN_B200_LIB; fi
  for cb in 8 4; do
  echo "== variant $v cb $cb"
  FBN_CB=$cb ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/exp_${v}_$cb.csv python tools/prof_run.py munin 672 0 1 0 > gpurun_out/exp_${v}_$cb.log 2>&1
  FBN_CB=$cb python tools/prof_run.py munin 1344 0 2 0 | tail -1
  done
done
