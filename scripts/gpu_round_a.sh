#!/bin/bash
# one gpurun call: C4 plans/s bench, then ncu launch lists + full captures for the C2 / C3 paths
set -x
python bench.py --workload c4 --steps 3 --warmup 1 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
python bench.py --workload c4 --steps 2 --warmup 1 --batch 4096 > gpurun_out/bench_c4_4096.json 2>> gpurun_out/bench_c4.err
C2="python bench.py --workload c2 --steps 2 --warmup 1 --no-cpu-baseline"
$C2 > gpurun_out/plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv $C2 > gpurun_out/ncu_c2.log 2>&1
$C2 > gpurun_out/plain_c2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fk_cull|narrow' -s 4 -c 4 -o gpurun_out/prof_c2 -f $C2 > gpurun_out/ncu_c2_full.log 2>&1
C3="python bench.py --workload c3 --batch 262144 --steps 2 --warmup 1 --no-cpu-baseline"
$C3 > gpurun_out/plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c3.csv $C3 > gpurun_out/ncu_c3.log 2>&1
cat gpurun_out/bench_c4.json gpurun_out/bench_c4_4096.json; tail -5 gpurun_out/bench_c4.err
