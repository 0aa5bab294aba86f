#!/usr/bin/env python
"""Summarise the ncu light-metrics pass of scripts/gpu_prof_r4.sh: per kernel name (B = 8 launches only for the
encoder kernels) count, mean duration and mean DRAM read + write bytes; writes the per-class traffic that bench.py
reports as roofline.traffic.  usage: summarize_launches.py gpurun_out/r4_launches.csv profiles/ [tag]"""
import collections
import csv
import json
import os
import re
import sys

src, outdir = sys.argv[1], sys.argv[2]
rows = []
with open(src) as fh:
    for line in fh:
        if line.startswith('"ID"'):
            break
    for r in csv.reader(fh):
        if len(r) >= 15:
            rows.append(r)
launch = collections.OrderedDict()
for r in rows:
    d = launch.setdefault(int(r[0]), {"name": r[4], "grid": r[8], "block": r[7]})
    d[r[12]] = float(r[14])
L = [launch[k] for k in sorted(launch)]


def short(n):
    n = re.sub(r"bsam::<unnamed>::", "", n)
    n = re.sub(r"^void ", "", n)
    return re.sub(r"\(.*$", "", n)[:90]


# encoder GEMMs of the 8-frame step: gemm_tc2_kernel<0> = qkv (and the global blocks' rel-pos G GEMM is gemm_tc_kernel<..,6,..>),
# <1> = lin1 + GELU, <2> = proj then lin2 alternately (both fp32-residual epilogues; proj has K = D, lin2 K = 4D: told
# apart by launch order)
cls = collections.defaultdict(list)
prev = ""
for d in L:
    n = short(d["name"])
    t = d.get("gpu__time_duration.sum", 0.0)
    by = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    key = None
    was, prev = prev, n
    if re.match(r"gemm_tc2_kernel<(\(int\))?0[,>]", n):
        key = "gemm_qkv"
    elif re.match(r"gemm_tc2_kernel<(\(int\))?1[,>]", n):
        key = "gemm_lin1"
    elif re.match(r"gemm_tc2_kernel<(\(int\))?[27][,>]", n):      # 2 = GM_F32, 7 = GM_F32_FOLD (same role with the folded LayerNorm)
        # fp32-residual epilogue: proj directly follows the block's attention kernel, lin2 follows lin1 (+GELU); the
        # remaining one per encode is the patch embedding
        key = "gemm_proj" if was.startswith("attn_") else "gemm_lin2" if was.startswith("gemm_tc2_kernel") else "gemm_patch"
    elif n.startswith("attn_window_tc_kernel"):
        key = "attn_window"
    elif n.startswith("attn_global_tc_kernel"):
        key = "attn_global"
    elif n.startswith("layernorm_bf16_kernel"):
        key = "layernorm"
    elif n.startswith("postprocess_kernel"):
        key = "postprocess"
    elif n.startswith("preprocess2_kernel") or n.startswith("preprocess_kernel"):
        key = "preprocess"
    if key:
        cls[key].append((t, by, d["grid"]))
byname = collections.defaultdict(list)
for d in L:
    byname[short(d["name"])].append((d.get("gpu__time_duration.sum", 0.0),
                                     d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)))
tag = sys.argv[3] if len(sys.argv) > 3 else "r4"
with open(os.path.join(outdir, f"{tag}_launch_summary.csv"), "w") as fh:
    fh.write("# ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none over "
             "`bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train` (all launches of the process, warm-up, latency and "
             "profiling passes included; cold-cache serialised timings)\nkernel,launches,mean_us,total_ms,mean_dram_MB\n")
    for n, v in sorted(byname.items(), key=lambda kv: -sum(x[0] for x in kv[1])):
        fh.write(f"\"{n}\",{len(v)},{sum(x[0] for x in v) / len(v) / 1e3:.2f},{sum(x[0] for x in v) / 1e6:.3f},"
                 f"{sum(x[1] for x in v) / len(v) / 1e6:.2f}\n")
traffic = {}
for k, v in cls.items():
    # the 8-frame launches are the long ones: keep launches within 2x of the class's longest
    tmax = max(x[0] for x in v)
    big = [x for x in v if x[0] > 0.5 * tmax]
    traffic[k] = sum(x[1] for x in big) / len(big)
traffic["_source"] = (f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none, capture {tag} "
                      "(scripts/gpu_prof_r4.sh): mean per launch over the 8-frame launches of each class")
json.dump(traffic, open(os.path.join(outdir, "ncu_traffic.json"), "w"), indent=1)
print(json.dumps(traffic, indent=1))
print("im2col launches:", sum(len(v) for n, v in byname.items() if "im2col" in n))
