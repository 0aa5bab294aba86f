#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_api.py -q -x > gpurun_out/r4c_train.log 2>&1; echo "rc=$?" >> gpurun_out/r4c_train.log
tail -8 gpurun_out/r4c_train.log
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --out gpurun_out/r4c_train_bench.json > gpurun_out/r4c_train_bench.log 2>&1; echo "rc=$?"
tail -5 gpurun_out/r4c_train_bench.log
