#!/bin/bash
# 256 x 128 pair tiles for small batches: model-level tests, B = 1 latency A/B (variant 2 = 1 auto / 2 = 256 only), B = 8 bench
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_path.py tests/test_gpu_api.py tests/test_gpu_fullsize.py tests/test_gpu_callers.py -m gpu -q -x > $O/bn128_tests.log 2>&1; echo "rc=$?" >> $O/bn128_tests.log
tail -n 4 $O/bn128_tests.log
timeout 600 python scripts/bench_latency_ab.py 2 1 2 4 40 > $O/bn128_latency.json 2> $O/bn128_latency.err; echo "lat rc=$?"
cat $O/bn128_latency.json
timeout 600 python bench.py --batch 1 --steps 20 --warmup 5 --no-cpu-baseline --no-train > $O/bn128_b1_bench.json 2> $O/bn128_b1_bench.err; echo "b1 rc=$?"
timeout 900 python bench.py --no-cpu-baseline --no-train > $O/bn128_b8_bench.json 2> $O/bn128_b8_bench.err; echo "b8 rc=$?"
python - <<'PY'
import json
for f in ("bn128_b1_bench", "bn128_b8_bench"):
    d = json.load(open(f"gpurun_out/{f}.json"))
    k = d["kernels"]
    print(f, round(d["value"], 2), round(d["ms_per_step"], 3), d["clocks"]["sm_mhz"],
          {n: round(k[n]["ms_per_step"], 3) for n in ("gemm_qkv", "gemm_proj", "gemm_lin1", "gemm_lin2")}, round(d["latency_b1_ms"]["device_ms"], 3))
PY
