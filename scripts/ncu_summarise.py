#!/usr/bin/env python
"""Turn the raw ncu CSV exports of scripts/gpu_prof.sh (gpurun_out/<tag>_*_raw.csv) into the committed summaries
under profiles/ and into profiles/ncu_traffic.json (DRAM bytes per launch per kernel class, read by bench.py for
`roofline.traffic`).  Usage: python scripts/ncu_summarise.py <tag>"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["ID", "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
# capture order of `cap gemm 'gemm_tc' 160 6` (bench.py --steps 1 --warmup 1): block 3 lin1, lin2, block 4 qkv, proj, ...
GEMM_ORDER = ["gemm_lin1", "gemm_lin2", "gemm_qkv", "gemm_proj", "gemm_lin1", "gemm_lin2"]


def main(tag):
    traffic = {}
    for name in ("gemm", "attn", "ln", "post", "dec", "aux"):
        path = os.path.join(ROOT, "gpurun_out", f"{tag}_{name}_raw.csv")
        if not os.path.exists(path):
            continue
        rows = list(csv.reader(open(path)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        idx = [hdr.index(w) for w in WANT if w in hdr]
        with open(os.path.join(ROOT, "profiles", f"{tag}_prof_{name}_summary.csv"), "w", newline="") as fh:
            w = csv.writer(fh)
            for r in rows:
                w.writerow([r[i] for i in idx])
        ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
        for n, r in enumerate(data):
            by = float(r[ir]) * UNIT[units[ir]] + float(r[iw]) * UNIT[units[iw]]
            k = r[ik]
            if name == "gemm":
                cls = GEMM_ORDER[n] if n < len(GEMM_ORDER) else None
            elif "attn_window" in k:
                cls = "attn_window"
            elif "attn_global" in k:
                cls = "attn_global"
            elif "layernorm" in k:
                cls = "layernorm"
            elif "postprocess" in k:
                cls = "postprocess"
            elif "preprocess" in k:
                cls = "preprocess"
            elif "mask_metrics" in k:
                cls = "mask_metrics"
            elif "rows_fc" in k:
                cls = "seg_project"
            else:
                cls = None
            if cls:
                traffic.setdefault(cls, []).append(by)
    out = {k: sum(v) / len(v) for k, v in traffic.items()}
    out["_source"] = f"ncu --set full --clock-control none, capture {tag}: dram__bytes_read.sum + dram__bytes_write.sum, mean per launch"
    with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1])
