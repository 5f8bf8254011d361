#!/bin/bash
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --workload c3 --steps 3 --no-cpu-baseline > gpurun_out/bench_c3_w3.json
python bench.py --workload c2 --steps 10 --no-cpu-baseline > gpurun_out/bench_c2_w3.json
python bench.py --workload c5 --steps 3 --no-cpu-baseline > gpurun_out/bench_c5_w3.json
python - <<PY
import json
for f in ["bench_c3_w3","bench_c2_w3","bench_c5_w3"]:
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"].get("kernel_ms_per_step"), d["stats_per_step"])
PY
