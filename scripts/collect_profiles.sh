#!/bin/bash
# copy one gpu_final.sh result set ($1 = directory) into profiles/ under the round's names
D=${1:-gpurun_out}; P=profiles
cp $D/bench_c2.json $P/r01_final_bench_c2.json; cp $D/bench_c3.json $P/r01_final_bench_c3.json; cp $D/bench_c5.json $P/r01_final_bench_c5.json
cp $D/bench_c4.json $P/r01_final_bench_c4_time0.5.json; cp $D/bench_c4_t025.json $P/r01_final_bench_c4_time0.25.json
for w in c2 c3 c4; do cp $D/bench_ref_$w.json $P/r01_final_bench_ref_$w.json; done
for w in c2 c3 c5; do cp $D/launches_$w.csv $P/r01_final_launches_$w.csv; done
cp $D/pytest_gpu.log $P/r01_final_pytest_gpu.log
for w in c2 c3 c5; do
  [ -f $D/prof_$w.ncu-rep ] || continue
  ncu -i $D/prof_$w.ncu-rep --page raw --csv > $P/r01_final_ncu_full_${w}_raw.csv 2>/dev/null
  ncu -i $D/prof_$w.ncu-rep --page details --csv > $P/r01_final_ncu_full_${w}_details.csv 2>/dev/null
done
