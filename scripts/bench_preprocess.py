#!/usr/bin/env python
"""Same-box A/B of the two preprocess kernel generations (debug variant 4: 0 = shipped, 1 = round-1 kernel):
8 raw 1080x1920 RGB u8 frames -> (8,3,1024,1024) bf16, CUDA events, algorithmic bytes / measured copy bandwidth."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

pkg = importlib.import_module("2handedafforder_b200")
lib = importlib.import_module("2handedafforder_b200._lib").load()
sam = pkg._build_sam(256, 2, 4, [1], max_batch=1, max_prompts=1).cuda()
peak = 6549.0
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
res = {}
for (H, W) in ((1080, 1920), (2160, 3840), (720, 1280)):
    B = 8
    raw = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device="cuda")
    nbytes = B * (H * W * 3 + 3 * 1024 * 1024 * 2)
    flush = torch.empty((256 << 20,), dtype=torch.uint8, device="cuda")
    for variant in (1, 0):
        assert lib.b200sam_debug_set_variant(4, variant) == 0
        for _ in range(3):
            sam.preprocess_frames(raw, torch.bfloat16)
        ts = []
        for _ in range(10):
            flush.zero_()                      # inputs out of L2
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sam.preprocess_frames(raw, torch.bfloat16)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        res[f"{H}x{W}:{'shipped' if variant == 0 else 'round1'}"] = {
            "ms": ms, "gbs": nbytes / ms / 1e6, "frac_hbm": nbytes / ms / 1e6 / peak}
lib.b200sam_debug_set_variant(4, 0)
print(json.dumps(res, indent=1))
if len(sys.argv) > 1:
    json.dump(res, open(sys.argv[1], "w"), indent=1)
