#!/bin/bash
set -x
mkdir -p gpurun_out
B200SAM_TEST_VARIANTS="5=1" timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r4r_tests_fold.log 2>&1; echo "rc=$?" >> gpurun_out/r4r_tests_fold.log
tail -25 gpurun_out/r4r_tests_fold.log | cut -c1-300
timeout 600 python scripts/bench_variant.py 5=0 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/r4r_bench_fold0.json 2>gpurun_out/r4r_fold0.err; echo rc=$?
timeout 600 python scripts/bench_variant.py 5=1 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/r4r_bench_fold1.json 2>gpurun_out/r4r_fold1.err; echo rc=$?
timeout 600 python scripts/bench_variant.py 5=0 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/r4r_bench_fold0b.json 2>gpurun_out/r4r_fold0.err; echo rc=$?
timeout 600 python scripts/bench_variant.py 5=1 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/r4r_bench_fold1b.json 2>gpurun_out/r4r_fold1.err; echo rc=$?
python - <<'PY'
import json
for f in ("fold0","fold1","fold0b","fold1b"):
    try:
        d=json.load(open(f"gpurun_out/r4r_bench_{f}.json")); print(f, round(d["value"],2), round(d["ms_per_step"],3), d["clocks"]["sm_mhz"], d["latency_b1_ms"]["device_ms"], {k:round(v["ms_per_step"],2) for k,v in d["kernels"].items() if not k.startswith("dec")})
    except Exception as e: print(f, "ERR", e)
PY
