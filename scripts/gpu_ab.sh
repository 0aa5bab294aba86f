#!/bin/bash
# Same-box A/B of two builds of the library (box-to-box clock / power-cap differences are larger than most kernel
# changes): runs the GPU tests on the new build, then bench.py twice each, alternating, and prints the per-class times.
# Usage: scripts/gpu_ab.sh <tag>   (baseline build at 2handedafforder_b200/libb200sam_base.so)
TAG=${1:-ab}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
for i in 1 2; do
  B200SAM_LIB=$PWD/2handedafforder_b200/libb200sam_base.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline \
      > $O/${TAG}_base_$i.json 2> $O/${TAG}_base_$i.err
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_new_$i.json 2> $O/${TAG}_new_$i.err
done
