#!/bin/bash
# same-box A/B of programmatic dependent launch (debug variant 6) on the current build, alternating three times
O=gpurun_out
mkdir -p $O
for r in 1 2 3; do
  for v in 0 1; do
    timeout 900 python scripts/bench_variant.py 6=$v -- --no-cpu-baseline --no-train > $O/pdl2_bench_${v}_$r.json 2> $O/pdl2_bench_${v}_$r.err; echo "bench 6=$v $r rc=$?"
  done
done
python - <<'PY'
import json
for v in (0, 1):
    for r in (1, 2, 3):
        d = json.load(open(f"gpurun_out/pdl2_bench_{v}_{r}.json"))
        print("6=%d" % v, r, round(d["value"], 2), round(d["ms_per_step"], 3), round(d["e2e"]["value"], 2), d["clocks"]["sm_mhz"],
              round(d["latency_b1_ms"]["device_ms"], 3))
PY
