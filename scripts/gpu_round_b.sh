#!/bin/bash
# C4 planner round anatomy: launch list of the first rounds (graph kernel nodes), plus time=1.0 budget runs
set -x
C4="python bench.py --workload c4 --steps 1 --warmup 1 --max-rounds 48"
$C4 > gpurun_out/plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -s 200 -c 700 --csv --log-file gpurun_out/launches_c4.csv $C4 > gpurun_out/ncu_c4.log 2>&1
tail -3 gpurun_out/ncu_c4.log
