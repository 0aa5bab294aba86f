#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_api.py tests/test_gpu_clip.py tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r4t_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r4t_tests.log
tail -3 gpurun_out/r4t_tests.log
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-train"
ncu --nvtx --nvtx-include "mask_decoder_right/" --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r4t_nvtx_decoder_right.csv $BENCH > gpurun_out/r4t_ncu.log 2>&1; echo "ncu rc=$?"
grep -c "gpu__time_duration" gpurun_out/r4t_nvtx_decoder_right.csv
grep "gpu__time_duration" gpurun_out/r4t_nvtx_decoder_right.csv | cut -d, -f5 | sort | uniq -c | sort -rn | head -12
