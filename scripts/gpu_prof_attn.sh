#!/bin/bash
# ncu --set full with source correlation of the two attention kernels inside a bench step (1 GPU).  The plain bench
# run must exit 0 first.  Usage: scripts/gpu_prof_attn.sh <tag>
TAG=${1:-r3}
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$BENCH > $O/${TAG}_plain.log 2>&1 || { echo "plain bench failed"; tail -5 $O/${TAG}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'attn_window_tc' --launch-skip 30 -c 2 -o $O/${TAG}_win -f \
    $BENCH > $O/${TAG}_ncu_win.log 2>&1; echo "ncu-win rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'attn_global_tc' --launch-skip 4 -c 1 -o $O/${TAG}_glb -f \
    $BENCH > $O/${TAG}_ncu_glb.log 2>&1; echo "ncu-glb rc=$?"
for k in win glb; do
  ncu -i $O/${TAG}_$k.ncu-rep --page raw --csv > $O/${TAG}_${k}_raw.csv 2>/dev/null
  ncu -i $O/${TAG}_$k.ncu-rep --page source --csv > $O/${TAG}_${k}_src.csv 2>/dev/null
done
ls -la $O
