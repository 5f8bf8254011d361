for v in d6 e6 e8 n4; do MPB_NO_NATIVE=1 MPB_LIB_PATH=build/variants/libmpb_$v.so python bench.py --also none --steps 5 --warmup 2 --no-cpu-baseline --no-latency --no-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'])"; done
