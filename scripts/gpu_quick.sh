#!/bin/bash
# quick check of a change: (optionally) GPU parity tests, then the three collision bench lines
if [ "$1" = "tests" ]; then python -m pytest tests -m gpu -x -q 2>&1 | tail -5; fi
for w in c2 c3 c5; do
  for F in ${INFLIGHT:-2}; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --inflight $F > gpurun_out/quick_$w.json 2> gpurun_out/quick_$w.err || tail -5 gpurun_out/quick_$w.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/quick_$w.json').read().strip().splitlines()[-1])
r=d['roofline']
print('$w F=$F', '%.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.3f'%d['ms_per_step'], r['kernel_ms_per_step'], 'fk', r['fk_table_kernel']['ms_per_step'], r['fk_table_kernel']['frac'], 'frac', r['frac'])
PY
  done
done
