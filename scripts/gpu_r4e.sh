#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -q > gpurun_out/r4e_train.log 2>&1; echo "rc=$?" >> gpurun_out/r4e_train.log
tail -12 gpurun_out/r4e_train.log
cat gpurun_out/train_parity.json
timeout 600 python scripts/prof_train.py --precision tf32 --out gpurun_out/r4e_prof_train_tf32.txt 2>&1 | tail -45
