#!/bin/bash
# gpurun_out/prof_<w>.ncu-rep -> profiles/r02_ncu_full_<w>_{raw,details}.csv (what bench.py's ncu_capture() reads)
for w in "$@"; do
  [ -f gpurun_out/prof_$w.ncu-rep ] || continue
  ncu -i gpurun_out/prof_$w.ncu-rep --page raw --csv > profiles/r02_ncu_full_${w}_raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_$w.ncu-rep --page details --csv > profiles/r02_ncu_full_${w}_details.csv 2>/dev/null
done
