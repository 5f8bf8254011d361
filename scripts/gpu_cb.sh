#!/bin/bash
# A/B of sphere_blk_kernel's register blocking (MPB_SPHERE_CB = 1: sphere_kernel, 2, 4) on the C3 workload
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for CB in 1 2 4; do
  MPB_SPHERE_CB=$CB python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/cb_$CB.json 2> gpurun_out/cb_$CB.err || tail -5 gpurun_out/cb_$CB.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/cb_$CB.json').read().strip().splitlines()[-1])
print('CB $CB c3 %.4g e2e %.4g'%(d['value'], d['e2e']['value']), d['roofline']['kernel_ms_per_step'], d['stats_per_step'])
PY
done
