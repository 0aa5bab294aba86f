#!/usr/bin/env python
"""Run bench.py in-process with kernel variant switches set first (process-wide, include/b200sam_debug.h):
    python scripts/bench_variant.py 5=1 -- --steps 10 --warmup 3 --no-cpu-baseline --no-train"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
args = sys.argv[1:]
split = args.index("--") if "--" in args else len(args)
lib = importlib.import_module("2handedafforder_b200._lib").load()
for kv in args[:split]:
    k, v = kv.split("=")
    assert lib.b200sam_debug_set_variant(int(k), int(v)) == 0
sys.argv = ["bench.py"] + args[split + 1:]
import bench  # noqa: E402

sys.exit(bench.main())
