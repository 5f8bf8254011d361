#!/bin/bash
# N-GPU bench lines (one process per GPU under torchrun, NCCL for the verdict gather): $1 = N
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for W in c2 c5; do
  $TR bench.py --gpus $N --workload $W --steps 10 --warmup 3 > gpurun_out/bench_${W}_n$N.json 2> gpurun_out/bench_${W}_n$N.err
  tail -1 gpurun_out/bench_${W}_n$N.json | cut -c1-220
done
$TR bench.py --gpus $N --workload c4 --steps 2 --warmup 1 > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err
tail -1 gpurun_out/bench_c4_n$N.json | cut -c1-220
for f in c2 c4; do tail -n 3 gpurun_out/bench_${f}_n$N.err; done
