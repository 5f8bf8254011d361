python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; grep -v "^  \|^$" gpurun_out/pytest_gpu.log | tail -40
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
def show(k,r):
    print(k, 'value %.4g e2e %.4g ms/step %.3f frac %.3f'%(r['value'], r['e2e']['value'], r['ms_per_step'], r['roofline']['frac'] or 0), r.get('parity'), r.get('latency'), r.get('cpu_baseline',{}).get('value'), r['roofline'].get('kernel_ms_per_step'), r.get('e2e'))
show('head', d)
for k,r in d.get('also',{}).items(): show(k,r)
PY
