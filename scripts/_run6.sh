CMD="python bench.py --workload c3 --also none --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-latency --batch 1048576"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'narrow2' -s 2 -c 2 -o gpurun_out/prof_n2 -f $CMD > gpurun_out/ncu_n2.log 2>&1
tail -2 gpurun_out/ncu_n2.log
