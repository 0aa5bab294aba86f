#!/bin/bash
# the GPU suite's encoder / path tests with the stand-alone LayerNorm (debug variant 5 = 1) and with PDL (6 = 1)
O=gpurun_out
mkdir -p $O
B200SAM_TEST_VARIANTS=5=1 timeout 1500 python -m pytest tests/test_gpu_ops.py tests/test_gpu_path.py tests/test_gpu_api.py tests/test_gpu_fullsize.py -m gpu -q -x > $O/var5_tests.log 2>&1; echo "rc=$?" >> $O/var5_tests.log
tail -3 $O/var5_tests.log
B200SAM_TEST_VARIANTS=6=1 timeout 1500 python -m pytest tests/test_gpu_ops.py tests/test_gpu_path.py tests/test_gpu_api.py tests/test_gpu_fullsize.py -m gpu -q -x > $O/var6_tests.log 2>&1; echo "rc=$?" >> $O/var6_tests.log
tail -3 $O/var6_tests.log
