#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py --batch 1 --steps 20 --warmup 5 --no-cpu-baseline --no-train > $O/b1_bench.json 2> $O/b1_bench.err; echo "rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/b1_bench.json"))
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
for k, v in d["kernels"].items():
    print(k, round(v["ms_per_step"], 3), v["launches_per_step"], v.get("tflops"))
PY
