#!/bin/bash
# eight GPUs: the driver's own scaling command for N = 8 (one bench line, gather verified)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/n8_bench.json 2> gpurun_out/n8_bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/n8_bench.json
