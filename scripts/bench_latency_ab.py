#!/usr/bin/env python
"""Same-process A/B of the B = 1 latency (bench.measure_latency_b1, the call sequence of LISA.evaluate) under a kernel
variant switch: python scripts/bench_latency_ab.py <which> <value_a> <value_b> [rounds] [iters]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

which, va, vb = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rounds = int(sys.argv[4]) if len(sys.argv) > 4 else 4
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 40
pkg = importlib.import_module("2handedafforder_b200")
lib = importlib.import_module("2handedafforder_b200._lib").load()
dev = torch.device("cuda", 0)
sam = pkg.build_sam_vit_h(None, max_batch=1, max_prompts=3)
g0 = torch.Generator().manual_seed(1)
with torch.no_grad():
    for name, p in sam.named_parameters():
        if "pos_embed" in name or "rel_pos" in name:
            p.copy_(0.02 * torch.randn(p.shape, generator=g0))
sam = sam.to(dev).to(torch.bfloat16)
sam.freeze_weights(True)
frame = torch.randn((1, 3, 1024, 1024), device=dev).to(torch.bfloat16)
seg = torch.randn((3, 256), device=dev).to(torch.bfloat16)
res = {str(va): [], str(vb): []}
for r in range(rounds):
    for v in (va, vb):
        assert lib.b200sam_debug_set_variant(which, v) == 0
        m = bench.measure_latency_b1(torch, sam, frame, seg, iters=iters)
        res[str(v)].append(round(m["device_ms"], 3))
lib.b200sam_debug_set_variant(which, 0)
print(json.dumps({"variant": which, "device_ms": res}))
