#!/bin/bash
# Round-2 final ncu pass (1 GPU): the plain bench first (must exit 0), then ONE light-metrics pass over a bench run
# (per-launch duration + DRAM bytes of every kernel: launch list, kernel shares and roofline.traffic come from it).
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train"
$BENCH > $O/r4_plain.json 2> $O/r4_plain.err || { echo "plain bench failed"; tail -5 $O/r4_plain.err; exit 1; }
cat $O/r4_plain.json | cut -c1-400
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
    --log-file $O/r4_launches.csv $BENCH > $O/r4_ncu_launch.log 2>&1; echo "ncu-list rc=$?"
wc -l $O/r4_launches.csv; tail -3 $O/r4_ncu_launch.log
