# throughput of every workload shape (tools/prof_run.py: <workload> <cases> <team> <reps> <mode>)
for a in "alarm 1000000 0 3 0" "hepar2 265216 0 2 0" "pigs 37888 0 2 0" "pigs 58944 0 2 2" "munin_small 3040 0 2 0" "munin 1344 0 2 0"; do
  timeout 600 python tools/prof_run.py $a 2>&1 | tail -1
done
echo "== FBN_NO_PERM=1"
for a in "pigs 37888 0 2 0" "pigs 58944 0 2 2" "munin 1344 0 2 0"; do
  FBN_NO_PERM=1 timeout 600 python tools/prof_run.py $a 2>&1 | tail -1
done
