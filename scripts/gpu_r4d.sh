#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -q -x > gpurun_out/r4d_train.log 2>&1; echo "rc=$?" >> gpurun_out/r4d_train.log
tail -25 gpurun_out/r4d_train.log
cat gpurun_out/train_parity.json
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --precision tf32 --out gpurun_out/r4d_train_bench_tf32.json 2>&1 | tail -3
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --precision fp32 --no-eager --cpu-sample 0 --out gpurun_out/r4d_train_bench_fp32.json 2>&1 | tail -3
