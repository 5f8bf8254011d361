#!/bin/bash
set -x
python -m pytest tests/test_gpu_planning.py -m gpu -x -q 2>&1 | tail -3
python bench.py --workload c4 --steps 3 --warmup 1 > gpurun_out/bench_c4_t1.json 2> gpurun_out/bench_c4.err
C4="python bench.py --workload c4 --steps 1 --warmup 1 --max-rounds 64"
$C4 > gpurun_out/plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -s 600 -c 600 --csv --log-file gpurun_out/launches_c4.csv $C4 > gpurun_out/ncu_c4.log 2>&1
python - <<PY
import json
for f in ["bench_c4_t1"]:
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["value"], d["e2e"]["value"], d.get("solve_rate"), d["ms_per_step"], d["roofline"].get("rounds_per_step"))
PY
tail -3 gpurun_out/bench_c4.err
