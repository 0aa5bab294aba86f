#!/bin/bash
mkdir -p gpurun_out
for r in 1 2 3; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$r bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/n2_run$r.json 2> gpurun_out/n2_run$r.err
python - <<PY
import json
d=json.load(open("gpurun_out/n2_run$r.json")); print("run$r", round(d["value"],1), round(d["e2e"]["value"],1), d["ms_per_step"], d["clocks"]["sm_mhz"])
PY
done
