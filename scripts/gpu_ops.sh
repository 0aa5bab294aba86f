#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -m gpu -q -x > $O/ops_tests.log 2>&1; echo "rc=$?" >> $O/ops_tests.log
tail -n 25 $O/ops_tests.log
