#!/bin/bash
# next-tile residual L2 prefetch: ops/path tests, same-box A/B against the previous commit's library, then the ncu capture
O=gpurun_out
mkdir -p $O
L=2handedafforder_b200/libb200sam.so
timeout 1800 python -m pytest tests/test_gpu_ops.py tests/test_gpu_path.py -m gpu -q -x > $O/r6b_tests.log 2>&1; echo "rc=$?" >> $O/r6b_tests.log
tail -3 $O/r6b_tests.log
cp $L build/libb200sam_new.so
for r in 1 2; do
  for w in new head; do
    cp build/libb200sam_$w.so $L
    timeout 900 python bench.py --no-cpu-baseline --no-train > $O/r6b_bench_${w}_$r.json 2> $O/r6b_bench_${w}_$r.err; echo "bench $w $r rc=$?"
  done
done
cp build/libb200sam_new.so $L
python - <<'PY'
import json
for w in ("new", "head"):
    for r in (1, 2):
        d = json.load(open(f"gpurun_out/r6b_bench_{w}_{r}.json"))
        k = d["kernels"]
        print(w, r, round(d["value"], 2), round(d["ms_per_step"], 3), round(d["e2e"]["value"], 2), d["clocks"]["sm_mhz"],
              {n: round(k[n]["ms_per_step"], 3) for n in ("gemm_qkv", "gemm_proj", "gemm_lin1", "gemm_lin2")},
              round(d["latency_b1_ms"]["device_ms"], 3))
PY
bash scripts/gpu_prof_r6.sh
