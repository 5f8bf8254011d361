CMD="python bench.py --workload c3 --also none --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-latency --batch 1048576"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sphere_grid|dense_kernel' -s 2 -c 2 -o gpurun_out/prof_sg -f $CMD > gpurun_out/ncu_sg.log 2>&1
tail -1 gpurun_out/ncu_sg.log
