#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: share of warp-stall samples per block of SASS instructions, with the
dominant opcodes and stall reasons of each block.  Usage: scripts/ncu_src_blocks.py <csv> [block=32] [min_pct=0.8]"""
import csv
import sys

path = sys.argv[1]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.8
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h, d = rows[hi], rows[hi + 1:]
tot = sum(int(r[2]) for r in d)
st = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not" not in x]
print("total samples", tot, "instructions", len(d))
for b in range(0, len(d), B):
    blk = d[b:b + B]
    s = sum(int(r[2]) for r in blk)
    if s < thr / 100 * tot:
        continue
    ops = {}
    for r in blk:
        t = r[1].strip().split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ops[op] = ops.get(op, 0) + 1
    agg = sorted(((sum(int(r[i] or 0) for r in blk), h[i][6:]) for i in st), reverse=True)[:3]
    n = max(1, sum(a for a, _ in agg))
    print(f"{b:5d} {100 * s / tot:5.1f}% exec {max(int(r[5]) for r in blk):>9d}",
          sorted(ops.items(), key=lambda kv: -kv[1])[:3], [(nm, round(100 * a / n)) for a, nm in agg])
