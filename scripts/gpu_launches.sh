#!/bin/bash
# ncu launch lists (device time per launch) of one short bench run per workload: $1 = workload list
for W in $1; do
  B=""; [ "$W" = "c3" ] && B="--batch 262144"
  CMD="python bench.py --workload $W $B --steps 2 --warmup 1 --no-cpu-baseline"
  $CMD > gpurun_out/plain_$W.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches_$W.csv $CMD > gpurun_out/ncu_$W.log 2>&1
done
