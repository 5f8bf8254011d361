#!/bin/bash
# One entry point for everything that runs under gpurun (outputs under gpurun_out/).
#   test [pytest args]   GPU test suite
#   bench [bench args]   one bench line (+ stderr)
#   quick                tests, then a short bench
#   launches <w> [args]  ncu launch list (gpu__time_duration) of a short single-workload run
#   full <w> [args]      ncu --set full capture of a short single-workload run (+ the FP32 op counters)
#   multi <N> [args]     bench under torchrun on N GPUs
mkdir -p gpurun_out
mode=$1; shift
case $mode in
  test)  python -m pytest tests -m gpu -q "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -25 gpurun_out/pytest_gpu.log ;;
  bench) python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err; cut -c1-3000 gpurun_out/bench.json ;;
  quick) python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -30 gpurun_out/pytest_gpu.log
         python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err; cut -c1-6000 gpurun_out/bench.json ;;
  launches) w=$1; shift
         CMD="python bench.py --workload $w --also none --steps 2 --warmup 1 --no-cpu-baseline --no-parity --no-latency $*"
         $CMD > gpurun_out/plain_$w.log 2>&1 &&
         ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$w.csv $CMD > gpurun_out/ncu_$w.log 2>&1
         echo "rc=$?"; tail -3 gpurun_out/ncu_$w.log ;;
  full)  w=$1; shift
         CMD="python bench.py --workload $w --also none --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-latency $*"
         $CMD > gpurun_out/plain_$w.log 2>&1 &&
         ncu --set full --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum \
             --clock-control none --import-source on -k regex:'fk_table|sphere|dense|narrow|expand|scatter|scan_bins|hits_compact' -s ${NCU_SKIP:-0} -c ${NCU_COUNT:-40} \
             -o gpurun_out/prof_$w -f $CMD > gpurun_out/ncu_full_$w.log 2>&1
         echo "rc=$?"; tail -3 gpurun_out/ncu_full_$w.log ;;
  multi) n=$1; shift
         python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n "$@" > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
         echo "rc=$?"; tail -5 gpurun_out/bench_n$n.err; cut -c1-3000 gpurun_out/bench_n$n.json ;;
esac
