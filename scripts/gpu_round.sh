#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list, ncu full captures.  Usage: scripts/gpu_round.sh <tag>
TAG=${1:-r1x}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
python bench.py --steps 10 --warmup 3 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err; echo "bench rc=$?"
cat $O/${TAG}_bench_n1.json
if [ "$2" != "noprof" ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu-list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'gemm_tc' --launch-skip 160 -c 8 -o $O/${TAG}_gemm -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_gemm.log 2>&1; echo "ncu-gemm rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'attn_(window|global)_tc' --launch-skip 34 -c 8 -o $O/${TAG}_attn -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_attn.log 2>&1; echo "ncu-attn rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'layernorm|postprocess|dec_' --launch-skip 90 -c 30 -o $O/${TAG}_mem -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu_mem.log 2>&1; echo "ncu-mem rc=$?"
fi
ls -la $O
