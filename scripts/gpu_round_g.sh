#!/bin/bash
set -x
C3="python bench.py --workload c3 --batch 262144 --steps 1 --warmup 1 --no-cpu-baseline"
$C3 > gpurun_out/plain_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fk_cull|narrow|expand' -s 5 -c 5 -o gpurun_out/prof_c3 -f $C3 > gpurun_out/ncu_c3_full.log 2>&1
tail -3 gpurun_out/ncu_c3_full.log
