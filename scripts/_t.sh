python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --also none --steps 5 --warmup 2 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['stats_per_step'], d['parity']['pair_mismatches'])"
python bench.py --workload c2 --also c5 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c2 value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['parity']['pair_mismatches']); d=d['also']['c5']; print('c5 value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['parity']['pair_mismatches'])"
