#!/usr/bin/env python
"""Print the same-box A/B of scripts/gpu_ab.sh: step time and per-class kernel times, baseline vs new build."""
import json
import sys

tag = sys.argv[1]
runs = {}
for arm in ("base", "new"):
    for i in (1, 2):
        try:
            runs[(arm, i)] = json.loads(open(f"gpurun_out/{tag}_{arm}_{i}.json").read().strip().splitlines()[-1])
        except Exception as e:  # noqa: BLE001
            print("missing", arm, i, e)
def avg(arm, f):
    v = [f(r) for (a, _), r in runs.items() if a == arm]
    return sum(v) / len(v) if v else float("nan")
print("clocks", [(k, r["clocks"]["sm_mhz"]) for k, r in runs.items()])
print(f"step ms   base {avg('base', lambda r: r['ms_per_step']):.2f}  new {avg('new', lambda r: r['ms_per_step']):.2f}")
print(f"frames/s  base {avg('base', lambda r: r['value']):.1f}  new {avg('new', lambda r: r['value']):.1f}")
ks = next(iter(runs.values()))["kernels"].keys()
for k in ks:
    b, n = avg("base", lambda r: r["kernels"][k]["ms_per_step"]), avg("new", lambda r: r["kernels"][k]["ms_per_step"])
    if abs(b - n) > 0.01 * max(b, 1e-9) and max(b, n) > 0.05:
        print(f"{k:16s} base {b:.3f}  new {n:.3f}  ({100 * (n / b - 1):+.1f}%)")
