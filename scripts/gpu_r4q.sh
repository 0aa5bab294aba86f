#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -q -x -k "layernorm_folded" > gpurun_out/r4q_fold_test.log 2>&1; echo "rc=$?" >> gpurun_out/r4q_fold_test.log
tail -15 gpurun_out/r4q_fold_test.log
timeout 600 python scripts/bench_variant.py 5=0 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/r4q_bench_fold0.json 2>gpurun_out/r4q_fold0.err; echo rc=$?
timeout 600 python scripts/bench_variant.py 5=1 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/r4q_bench_fold1.json 2>gpurun_out/r4q_fold1.err; echo rc=$?
python - <<'PY'
import json
for f in ("gpurun_out/r4q_bench_fold0.json","gpurun_out/r4q_bench_fold1.json"):
    try:
        d=json.load(open(f)); print(f, d["value"], d["ms_per_step"], {k:round(v["ms_per_step"],3) for k,v in d["kernels"].items() if not k.startswith("dec")})
    except Exception as e: print(f, "ERR", e)
PY
tail -3 gpurun_out/r4q_fold1.err
