# Application-level run of the CLI drop-in on 1M ALARM cases (-f3 / -f4 ingestion and the pt writer):
# run 1 writes the posteriors as a potential-table file (-o), run 2 reads that 1 GB file back as the -f4 golden and scores.
set -e
N=${1:-1000000}
D=build/app
mkdir -p $D
cp tests/golden/alarm/alarm.xml $D/
python - <<PY
import numpy as np, time, sys
sys.path.insert(0, ".")
from fastbn_b200 import synth
t = np.load("tests/golden/alarm_tree_ref.npz")
t0 = time.time()
off, ev, ew = synth.uniform_evidence(t["var_dims"], $N, 7, seed=20260101)
synth.write_libsvm("$D/cases", off, ev, ew)
synth.write_empty_pt("$D/empty_pt")
print("generated %d cases in %.1f s" % ($N, time.time() - t0))
PY
ls -la $D/cases
export FASTBN_TIMING=1
echo "== run 1 (-f4 empty, -o out_pt)"
./fastbn_b200/bin/BayesianNetwork -D $D -f0 alarm.xml -f3 cases -f4 empty_pt -o out_pt | tail -4
ls -la $D/out_pt
echo "== run 2 (-f4 out_pt)"
./fastbn_b200/bin/BayesianNetwork -D $D -f0 alarm.xml -f3 cases -f4 out_pt | tail -4
echo "== run 3 (2 GPUs if present)"
./fastbn_b200/bin/BayesianNetwork -D $D -f0 alarm.xml -f3 cases -f4 out_pt -G ${GPUS:-1} | tail -4
rm -f $D/out_pt $D/cases
