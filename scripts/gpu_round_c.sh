#!/bin/bash
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --workload c4 --steps 3 --warmup 1 > gpurun_out/bench_c4_t1.json 2> gpurun_out/bench_c4.err
python bench.py --workload c4 --steps 2 --warmup 1 --plan-time 60 > gpurun_out/bench_c4_t60.json 2>> gpurun_out/bench_c4.err
python bench.py --workload c3 --steps 3 --no-cpu-baseline > gpurun_out/bench_c3_w2.json
python - <<PY
import json
for f in ["bench_c4_t1","bench_c4_t60","bench_c3_w2"]:
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["value"], d["e2e"]["value"], d.get("solve_rate"), d["ms_per_step"], d["roofline"].get("rounds_per_step"), d["roofline"].get("kernel_ms_per_step"))
PY
tail -3 gpurun_out/bench_c4.err
