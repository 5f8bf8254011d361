#!/bin/bash
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --workload c4 --steps 3 --warmup 1 > gpurun_out/bench_c4_t1.json 2> gpurun_out/bench_c4.err
python bench.py --workload c4 --steps 3 --warmup 1 --plan-time 0.25 > gpurun_out/bench_c4_t025.json 2>> gpurun_out/bench_c4.err
C3="python bench.py --workload c3 --batch 262144 --steps 2 --warmup 1 --no-cpu-baseline"
$C3 > gpurun_out/plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c3.csv $C3 > gpurun_out/ncu_c3.log 2>&1
python - <<PY
import json
for f in ["bench_c4_t1","bench_c4_t025"]:
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["value"], d["e2e"]["value"], d.get("solve_rate"), d["ms_per_step"], d["roofline"].get("rounds_per_step"))
PY
tail -3 gpurun_out/bench_c4.err
