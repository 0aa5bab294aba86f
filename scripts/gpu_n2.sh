#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_clip.py -q -m gpu > gpurun_out/n2_clip_tests.log 2>&1; echo "rc=$?" >> gpurun_out/n2_clip_tests.log
tail -4 gpurun_out/n2_clip_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo "bench rc=$?"
tail -c 1200 gpurun_out/n2_bench.json
