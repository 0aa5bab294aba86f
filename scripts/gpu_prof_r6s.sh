#!/bin/bash
# ncu --set full with the source page of the proj launch (second of the captured pair GEMMs): where its warps wait
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train"
ncu --set full --clock-control none --import-source on -k regex:'gemm_tc2' --launch-skip 400 -c 2 -o $O/r6s_gemm -f $BENCH > $O/r6s_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/r6s_gemm.ncu-rep --page source --csv --kernel-id :::2 > $O/r6s_source_proj.csv 2>/dev/null
ncu -i $O/r6s_gemm.ncu-rep --page raw --csv > $O/r6s_raw.csv 2>/dev/null
rm -f $O/r6s_gemm.ncu-rep
ls -la $O/r6s_*
