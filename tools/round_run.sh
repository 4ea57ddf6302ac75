#!/bin/bash
# Round-end style run on one B200 (tag = $1): GPU tests, smoke, ncu launch lists of every workload -> profiles/traffic.json,
# one --set full capture of the headline kernel (shared-memory traffic), then the bench line of all five configs and the
# reference arm (plain runs, after the captures so that the line carries traffic figures of this kernel revision).
cd "$(dirname "$0")/.."
tag=${1:-rN}
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -1 gpurun_out/${tag}_smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_gpu_tests.log 2>&1; tail -1 gpurun_out/${tag}_gpu_tests.log
bash tools/r2_traffic.sh $tag
python tools/prof_run.py alarm 400000 0 2 > gpurun_out/${tag}_plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:jt_resident -s 1 -c 1 -f -o gpurun_out/${tag}_ncu_full_resident_alarm python tools/prof_run.py alarm 400000 0 2 > gpurun_out/${tag}_ncu_full.log 2>&1
team=$(grep -o "team [0-9]*" gpurun_out/${tag}_plain_full.log | head -1 | cut -d' ' -f2)
python tools/update_traffic.py --smem alarm gpurun_out/${tag}_ncu_full_resident_alarm.ncu-rep 400000 $team && cp profiles/traffic.json gpurun_out/${tag}_traffic.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_bench_ref.err
python bench.py > gpurun_out/${tag}_bench_all_configs.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<P
import json
d=json.load(open("gpurun_out/${tag}_bench_all_configs.json"))
print("ALARM", d["value"], "e2e", d["e2e"]["value"], "full posteriors", d["e2e"]["full_posteriors"]["value"], d["roofline"]["frac"], (d["roofline"].get("on_chip") or {}).get("frac"), d["parity_prefix"]["ok"])
for k,v in d["workloads"].items(): print(k, v["value"], v["e2e"]["value"], v["dram_frac"], v["parity_prefix"]["ok"])
P
