for v in 1 4 8 12 16; do MPB_REFILL_MIN=$v python bench.py --also none --steps 5 --warmup 2 --no-cpu-baseline --no-latency --no-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('refill_min=$v', 'value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'])"; done
MPB_REFILL_MIN=8 python bench.py --workload c2 --also c5 --steps 10 --warmup 3 --no-cpu-baseline --no-latency --no-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c2 value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step']); d=d['also']['c5']; print('c5 value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'])"
