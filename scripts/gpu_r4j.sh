#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_callers.py -q -x -m gpu > gpurun_out/r4j_callers.log 2>&1; echo "rc=$?" >> gpurun_out/r4j_callers.log
tail -12 gpurun_out/r4j_callers.log
timeout 600 python scripts/bench_preprocess.py gpurun_out/r4j_preprocess_ab.json 2>&1 | tail -30
