#!/bin/bash
# light-metrics ncu pass over one bench run of the round's last build: launch list, kernel shares, roofline.traffic
O=gpurun_out
mkdir -p $O
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train"
timeout 200 $BENCH > $O/r6_plain.json 2> $O/r6_plain.err || { echo "plain bench failed"; tail -n 5 $O/r6_plain.err; exit 1; }
timeout 260 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
    --log-file $O/r6_launches.csv $BENCH > $O/r6_ncu_launch.log 2>&1; echo "ncu-list rc=$?"
wc -l $O/r6_launches.csv; tail -n 2 $O/r6_ncu_launch.log | cut -c1-300
