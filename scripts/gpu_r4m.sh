#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -q > gpurun_out/r4m_train.log 2>&1; echo "rc=$?" >> gpurun_out/r4m_train.log
tail -8 gpurun_out/r4m_train.log
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --precision tf32 --no-eager --cpu-sample 0 --out gpurun_out/r4m_train_bench_tf32.json 2>&1 | tail -2
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --precision fp32 --no-eager --cpu-sample 0 --out gpurun_out/r4m_train_bench_fp32.json 2>&1 | tail -2
timeout 600 python scripts/prof_train.py --precision tf32 --out gpurun_out/r4m_prof_train_tf32.txt 2>&1 | grep -v "at::native" | head -24
