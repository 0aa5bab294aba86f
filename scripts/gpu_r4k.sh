#!/bin/bash
set -x
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:preprocess2_kernel --launch-skip 3 -c 1 -o gpurun_out/r4k_pre2 -f python scripts/bench_preprocess.py > gpurun_out/r4k_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r4k_pre2.ncu-rep --page raw --csv > gpurun_out/r4k_pre2_raw.csv 2>/dev/null
ncu -i gpurun_out/r4k_pre2.ncu-rep --page source --csv > gpurun_out/r4k_pre2_source.csv 2>/dev/null
ls -la gpurun_out | tail -5
