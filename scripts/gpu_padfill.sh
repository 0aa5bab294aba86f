#!/bin/bash
# window-pad rows written by the previous block's lin2 launch (variant 7 = 0) against the fill kernel (7 = 1)
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_path.py tests/test_gpu_api.py tests/test_gpu_fullsize.py -m gpu -q -x > $O/padfill_tests.log 2>&1; echo "rc=$?" >> $O/padfill_tests.log
tail -n 4 $O/padfill_tests.log
for r in 1 2 3; do
  for v in 0 1; do
    timeout 600 python scripts/bench_variant.py 7=$v -- --no-cpu-baseline --no-train > $O/padfill_bench_${v}_$r.json 2> $O/padfill_bench_${v}_$r.err; echo "bench 7=$v $r rc=$?"
  done
done
python - <<'PY'
import json
for v in (0, 1):
    for r in (1, 2, 3):
        d = json.load(open(f"gpurun_out/padfill_bench_{v}_{r}.json"))
        print("7=%d" % v, r, round(d["value"], 2), round(d["ms_per_step"], 3), round(d["e2e"]["value"], 2), d["clocks"]["sm_mhz"],
              d["gpu_launches"], round(d["kernels"]["gemm_qkv"]["ms_per_step"], 3), round(d["kernels"]["gemm_lin2"]["ms_per_step"], 3),
              round(d["latency_b1_ms"]["device_ms"], 3))
PY
