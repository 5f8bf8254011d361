#!/bin/bash
# gpurun_out/ -> profiles/r02_* (run here after scripts/gpu_evidence.sh came back)
D=gpurun_out; P=profiles; T=${1:-r02}
cp $D/pytest_gpu.log $P/${T}_pytest_gpu.log
grep '^{' $D/bench.json | tail -1 > $P/${T}_bench.json
grep '^{' $D/bench_ref.json | tail -1 > $P/${T}_bench_ref.json
for n in 2 4 8; do [ -f $D/bench_n$n.json ] && grep '^{' $D/bench_n$n.json | tail -1 > $P/${T}_bench_n$n.json; done
for w in c2 c3 c5; do
  [ -f $D/launches_$w.csv ] && cp $D/launches_$w.csv $P/${T}_launches_$w.csv
  [ -f $D/prof_$w.ncu-rep ] || continue
  ncu -i $D/prof_$w.ncu-rep --page raw --csv > $P/${T}_ncu_full_${w}_raw.csv 2>/dev/null
  ncu -i $D/prof_$w.ncu-rep --page details --csv > $P/${T}_ncu_full_${w}_details.csv 2>/dev/null
done
ls -la $P | grep ${T}_ | wc -l
