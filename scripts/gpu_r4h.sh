#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r4h_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r4h_noalloc_alone.log
tail -15 gpurun_out/r4h_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r4h_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r4h_tests.log
tail -15 gpurun_out/r4h_tests.log
cat gpurun_out/train_parity.json
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --precision tf32 --out gpurun_out/r4h_train_bench_tf32.json 2>&1 | tail -2
timeout 900 python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --precision fp32 --no-eager --cpu-sample 0 --out gpurun_out/r4h_train_bench_fp32.json 2>&1 | tail -2
timeout 600 python scripts/prof_train.py --precision tf32 --out gpurun_out/r4h_prof_train_tf32.txt 2>&1 | grep -v "at::native" | tail -30
