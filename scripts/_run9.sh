for cfgs in "1024 1024 1024" "4096 1024 1024" "4096 2048 1024" "4096 1024 512" "4096 2048 2048" "8192 4096 1024"; do set -- $cfgs
python bench.py --workload c4 --also none --batch $1 --plan-slots $2 --plan-rounds-per-query $3 --steps 2 --warmup 1 --no-cpu-baseline --no-parity > gpurun_out/c4.out 2>gpurun_out/c4.err; echo "Q=$1 slots=$2 R=$3 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/c4.out')); print('   plans/s %.0f e2e %.0f solve %.3f ms/step %.1f'%(d['value'], d['e2e']['value'], d['solve_rate'], d['ms_per_step']), d['budget']['grow_calls_per_query'], d['roofline']['rounds_per_step'])" || tail -4 gpurun_out/c4.err; done
