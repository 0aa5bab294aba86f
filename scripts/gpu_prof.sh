#!/bin/bash
# ncu pass only (bench.py must already have exited 0 without ncu in this round): launch list + full captures,
# exported to CSV on the box so the summaries survive the 64 MiB gpurun_out cap.  Usage: scripts/gpu_prof.sh <tag>
TAG=${1:-r1x}
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu-list rc=$?"
cap() {  # name regex skip count
  ncu --set full --clock-control none --import-source on -k regex:"$2" --launch-skip $3 -c $4 -o $O/${TAG}_$1 -f \
      $BENCH > $O/${TAG}_ncu_$1.log 2>&1; echo "ncu-$1 rc=$?"
  ncu -i $O/${TAG}_$1.ncu-rep --page raw --csv > $O/${TAG}_$1_raw.csv 2>/dev/null
}
cap gemm 'gemm_tc' 160 6
cap attn 'attn_(window|global)_tc' 34 6
cap ln 'layernorm' 70 2
cap post 'postprocess' 2 2
cap dec 'dec_' 22 22
rm -f $O/${TAG}_dec.ncu-rep
cap aux 'preprocess_kernel|rows_fc_kernel|mask_metrics_kernel|seg_scan_kernel' 0 10
# gpurun copies back at most 64 MiB: the CSV exports are what gets committed, drop the largest reports until we fit
while [ $(du -sm $O | cut -f1) -ge 48 ]; do
  big=$(ls -S $O/*.ncu-rep 2>/dev/null | head -1); [ -z "$big" ] && break; rm -f "$big"; echo "dropped $big"
done
du -sh $O; ls -la $O
