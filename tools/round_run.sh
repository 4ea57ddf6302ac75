# round-end style run on one B200: gpu tests, smoke, bench lines of every workload, ncu launch list + full capture of the bench command
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -1 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_alarm.json 2> gpurun_out/bench_alarm.err
for w in hepar2 hepar2_50 pigs munin_small munin; do
  python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1c_launches_bench_alarm.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:jt_fused -s 3 -c 1 -o gpurun_out/r1c_fused_alarm -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:jt_fused -s 1 -c 1 -o gpurun_out/r1c_fused_pigs -f python tools/prof_run.py pigs 4736 0 2 0 > gpurun_out/ncu_full_pigs.log 2>&1
