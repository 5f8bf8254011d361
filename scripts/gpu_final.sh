#!/bin/bash
# final evidence run of the round: GPU tests, bench lines (both arms), launch lists, one full ncu capture per path
set -x
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --workload c3 --steps 5 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python bench.py --workload c5 --steps 5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err
python bench.py --workload c4 --steps 4 --warmup 2 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
python bench.py --workload c4 --steps 4 --warmup 2 --plan-time 0.25 > gpurun_out/bench_c4_t025.json 2>> gpurun_out/bench_c4.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c2.json 2>&1
python bench.py --impl reference --workload c4 --steps 2 --warmup 0 > gpurun_out/bench_ref_c4.json 2>&1
python bench.py --impl reference --workload c3 --steps 2 --warmup 1 > gpurun_out/bench_ref_c3.json 2>&1
bash scripts/gpu_launches.sh "c2 c3 c5"
C2="python bench.py --workload c2 --steps 1 --warmup 1 --no-cpu-baseline"
$C2 > gpurun_out/plain_c2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fk_table|sphere|dense|narrow|scatter' -s 6 -c 6 -o gpurun_out/prof_c2 -f $C2 > gpurun_out/ncu_c2_full.log 2>&1
C3="python bench.py --workload c3 --batch 262144 --steps 1 --warmup 1 --no-cpu-baseline"
$C3 > gpurun_out/plain_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fk_table|sphere|dense|narrow|expand|scatter' -s 10 -c 10 -o gpurun_out/prof_c3 -f $C3 > gpurun_out/ncu_c3_full.log 2>&1
C5="python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu-baseline"
$C5 > gpurun_out/plain_c5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fk_table|sphere|dense|narrow|scatter' -s 6 -c 6 -o gpurun_out/prof_c5 -f $C5 > gpurun_out/ncu_c5_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log
