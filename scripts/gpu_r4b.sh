#!/bin/bash
# training row: primitives + whole step
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -q -x > gpurun_out/r4b_train.log 2>&1; echo "rc=$?" >> gpurun_out/r4b_train.log
tail -40 gpurun_out/r4b_train.log
timeout 600 python -m pytest tests/test_gpu_api.py -q -x > gpurun_out/r4b_api.log 2>&1; echo "rc=$?" >> gpurun_out/r4b_api.log
tail -15 gpurun_out/r4b_api.log
