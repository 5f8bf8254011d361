python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "expand_variants or env_full" 2>&1 | tail -3
python bench.py --also none --steps 5 --warmup 2 --no-cpu-baseline --no-latency --no-parity 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value %.4g ms %.3f'%(d['value'], d['ms_per_step']), d['roofline']['kernel_ms_per_step'], d['stats_per_step'])"
CMD="python bench.py --workload c3 --also none --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-latency --batch 262144"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'expand_lane' -s 2 -c 1 -o gpurun_out/prof_expl -f $CMD > gpurun_out/ncu_expl.log 2>&1
tail -2 gpurun_out/ncu_expl.log
