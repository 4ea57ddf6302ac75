# throughput of every workload shape (tools/prof_run.py: <workload> <cases> <team> <reps> <mode>)
for v in ${VARIANTS:-default}; do
  if [ "$v" != default ]; then export FASTBN_B200_LIB=$PWD/build/var_$v/libfastbn_b200.so; else unset FASTBN_B200_LIB; fi
  echo "== variant $v"
  for a in "alarm 1000000 0 3 0" "hepar2 265216 0 2 0" "pigs 37888 0 2 0" "pigs 58944 0 2 2" "munin_small 3040 0 2 0" "munin 1344 0 2 0"; do
    timeout 600 python tools/prof_run.py $a 2>&1 | tail -1
  done
done
