#!/bin/bash
# round-2 late check: implicit convolutions
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -q -x -k "tma_gathered" > gpurun_out/r4a_new.log 2>&1; echo "rc=$?" >> gpurun_out/r4a_new.log
tail -5 gpurun_out/r4a_new.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r4a_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r4a_tests.log
tail -5 gpurun_out/r4a_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r4a_bench.json 2> gpurun_out/r4a_bench.err; echo "rc=$?"
cat gpurun_out/r4a_bench.json
