T0=$(date +%s)
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$? wall $(( $(date +%s) - T0 )) s"; tail -3 gpurun_out/bench.err
T0=$(date +%s)
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$? wall $(( $(date +%s) - T0 )) s"; tail -3 gpurun_out/bench_ref.err
python - <<'PY'
import json
for f in ("gpurun_out/bench.json","gpurun_out/bench_ref.json"):
    d=json.load(open(f))
    def show(k,r):
        print(k, '%s value %.4g e2e %.4g ms/step %.3f'%(r['metric'][:20], r['value'], r['e2e']['value'], r['ms_per_step']), 'frac', (r.get('roofline') or {}).get('frac'), 'cpu', (r.get('cpu_baseline') or {}).get('value'), (r.get('cpu_baseline') or {}).get('cores'), r.get('solve_rate'), r.get('budget'))
    show('head', d)
    for k,r in d.get('also',{}).items(): show(k,r)
    if 'latency' in d: print(d['latency'])
PY
