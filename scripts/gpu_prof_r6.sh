#!/bin/bash
# ncu --set full (with source) of five consecutive pair GEMMs after the branch-free epilogues (lin2 / qkv / proj / lin1 ...)
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train"
ncu --set full --clock-control none --import-source on -k regex:'gemm_tc2' --launch-skip 400 -c 5 -o $O/r6d_gemm -f $BENCH > $O/r6_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/r6d_gemm.ncu-rep --page raw --csv > $O/r6d_gemm_raw.csv 2>/dev/null
ls -la $O/r6_*
