# ALARM throughput of flat-kernel builds with different batch sizes (build/var_b<C>_<P>/)
for v in ${VARIANTS:-default}; do
  if [ "$v" != default ]; then export FASTBN_B200_LIB=$PWD/build/var_$v/libfastbn_b200.so; else unset FASTBN_B200_LIB; fi
  echo "== variant $v"; python tools/prof_run.py alarm 1000000 0 3 0 | tail -1
done
