#!/bin/bash
# Evidence set of the round on ONE GPU: GPU tests, the default bench line of both arms, ncu launch lists and full captures
# (each ncu pass after the same command exited 0 without it).  Outputs under gpurun_out/; scripts/collect_evidence.sh copies
# the summaries into profiles/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
for w in c3 c2 c5; do
  B=""; [ "$w" = "c3" ] && B="--batch 1048576"; [ "$w" = "c5" ] && B="--batch 1048576"
  bash scripts/gpu_run.sh launches $w $B > /dev/null 2>&1
  # exactly the second pass of 2^20 states (the first belongs to the calibration call): the -k filter of gpu_run.sh matches
  # 12 kernels per pass with BVH objects (c3), 7 without
  K=7; [ "$w" = "c3" ] && K=12
  NCU_SKIP=$K NCU_COUNT=$K bash scripts/gpu_run.sh full $w $B | tail -1
done
