#!/bin/bash
set -x
mkdir -p gpurun_out
B200SAM_TEST_VARIANTS="5=1" timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r4s_tests_standalone_ln.log 2>&1; echo "rc=$?" >> gpurun_out/r4s_tests_standalone_ln.log
tail -4 gpurun_out/r4s_tests_standalone_ln.log
bash scripts/gpu_final.sh
