python -m pytest tests/test_gpu_parity.py tests/test_gpu_async.py -m gpu -q -x 2>&1 | tail -3
for v in 0 1; do MPB_DENSE_V1=$v python bench.py --also none --steps 5 --warmup 2 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('dense_v1=$v value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['stats_per_step']['support_evals'], d['parity']['pair_mismatches'])"; done
for v in 0 1; do MPB_DENSE_V1=$v python bench.py --workload c2 --also c5 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('v1=$v c2 value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['parity']['pair_mismatches']); d=d['also']['c5']; print('c5 value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['parity']['pair_mismatches'])"; done
