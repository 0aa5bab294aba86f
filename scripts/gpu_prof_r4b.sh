#!/bin/bash
# ncu --set full of the GEMMs of one transformer block with the folded LayerNorm (lin2 -> qkv -> ... -> proj -> lin1 -> lin2)
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train"
ncu --set full --clock-control none --import-source on -k regex:'gemm_tc2' --launch-skip 400 -c 5 -o $O/r4v_gemm -f $BENCH > $O/r4v_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/r4v_gemm.ncu-rep --page raw --csv > $O/r4v_gemm_raw.csv 2>/dev/null
ls -la $O/r4v_gemm_raw.csv
rm -f $O/r4v_gemm.ncu-rep
