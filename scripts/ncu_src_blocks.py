#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export (SASS view): share of warp-stall samples per block of SASS
instructions, with the dominant opcodes and stall reasons of each block, for one kernel section of the file.
Usage: scripts/ncu_src_blocks.py <csv> [block=32] [min_pct=0.8] [section=0]"""
import csv
import sys

path = sys.argv[1]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.8
sec = int(sys.argv[4]) if len(sys.argv) > 4 else 0
rows = list(csv.reader(open(path)))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hi = heads[sec]
end = heads[sec + 1] - 1 if sec + 1 < len(heads) else len(rows)
h = rows[hi]
d = [r for r in rows[hi + 1:end] if len(r) == len(h) and r[0].startswith("0x")]
tot = sum(int(r[2] or 0) for r in d)
st = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not" not in x]
KEY = ("LDTM", "STTM", "MUFU", "SYNCS", "UTCHMMA", "UTMALDG", "BAR", "LDS", "STS", "STG", "LDG", "WARPSYNC", "UTCBAR",
       "FENCE", "ELECT", "NANOSLEEP")
print("kernel", rows[hi - 1][1][:100] if hi else "")
print("total samples", tot, "instructions", len(d))
for b in range(0, len(d), B):
    blk = d[b:b + B]
    s = sum(int(r[2] or 0) for r in blk)
    if s < thr / 100 * tot:
        continue
    ops = {}
    for r in blk:
        t = r[1].strip().split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ops[op] = ops.get(op, 0) + 1
    agg = sorted(((sum(int(r[i] or 0) for r in blk), h[i][6:]) for i in st), reverse=True)[:3]
    n = max(1, sum(a for a, _ in agg))
    keys = {k: v for k, v in ops.items() if k.startswith(KEY)}
    top = sorted(ops.items(), key=lambda kv: -kv[1])[:3]
    hot = max(blk, key=lambda r: int(r[2] or 0))
    print(f"{b:5d} {100 * s / tot:5.1f}% exec {max(int(r[5] or 0) for r in blk):>9d}", top, keys,
          [(nm, round(100 * a / n)) for a, nm in agg], "| hot:", hot[1].strip()[:48], hot[2])
