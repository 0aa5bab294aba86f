#!/bin/bash
# Round-3 ncu pass (1 GPU): launch list of one bench step + full captures of the two attention kernels and one of each
# big GEMM; raw-page CSVs are exported on the box.  The plain bench run must exit 0 first.  Usage: scripts/gpu_prof_r3.sh <tag>
TAG=${1:-r3}
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$BENCH > $O/${TAG}_plain.json 2> $O/${TAG}_plain.err || { echo "plain bench failed"; tail -5 $O/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/${TAG}_launches.csv \
    $BENCH > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu-list rc=$?"
cap() {  # name regex skip count
  ncu --set full --clock-control none --import-source on -k regex:"$2" --launch-skip $3 -c $4 -o $O/${TAG}_$1 -f \
      $BENCH > $O/${TAG}_ncu_$1.log 2>&1; echo "ncu-$1 rc=$?"
  ncu -i $O/${TAG}_$1.ncu-rep --page raw --csv > $O/${TAG}_$1_raw.csv 2>/dev/null
}
cap win 'attn_window_tc' 30 2
cap glb 'attn_global_tc' 4 1
cap gemm 'gemm_tc2' 200 4
cap post 'postprocess_kernel' 4 2
rm -f $O/${TAG}_gemm.ncu-rep $O/${TAG}_post.ncu-rep
ls -la $O | tail -20
