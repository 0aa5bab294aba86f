#!/usr/bin/env python
"""Same-box A/B of the attention kernel variants (include/b200sam_debug.h): times the windowed (S = 14) and global
(S = 64) attention op at the shapes of one ViT-H encoder block for a batch of 8 frames, every variant on the same
inputs in one process, and checks each variant's output against the first one listed (window variants: 0 = shipped,
9 = round-1 kernel, 4 = shipped + L2 prefetch).
Usage: python scripts/bench_attn.py [--win 0,1] [--glb 0,1] [--iters 20] [--batch 8]"""
import argparse
import ctypes as C
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

_lib = importlib.import_module("2handedafforder_b200._lib")


def run(lib, S, nb, heads, hd, variants, iters, which):
    g = torch.Generator().manual_seed(30 + S)
    rows = nb * S * S
    qkv = (torch.randn((rows, 3 * heads * hd), generator=g) * 1.5).to(torch.bfloat16).cuda()
    rel_h = (torch.randn((2 * S - 1, hd), generator=g) * 0.3).cuda()
    rel_w = (torch.randn((2 * S - 1, hd), generator=g) * 0.3).cuda()
    ref = None
    res = {}
    for v in variants:
        assert lib.b200sam_debug_set_variant(which, v) == 0
        out = torch.zeros((rows, heads * hd), dtype=torch.bfloat16, device="cuda")
        ms = C.c_float()
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        rc = lib.b200sam_debug_attention_bench(C.c_void_p(qkv.data_ptr()), nb, S, heads, hd, C.c_void_p(rel_h.data_ptr()),
                                               C.c_void_p(rel_w.data_ptr()), C.c_void_p(out.data_ptr()), iters,
                                               C.byref(ms), st)
        torch.cuda.synchronize()
        entry = {"rc": rc, "ms": ms.value}
        if rc == 0:
            if ref is None:
                ref = out.float()
            d = (out.float() - ref).abs()
            entry["max_abs_diff_vs_v%d" % variants[0]] = float(d.max())
            entry["finite"] = bool(torch.isfinite(out.float()).all())
        res[f"v{v}"] = entry
    lib.b200sam_debug_set_variant(which, 0)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--win", default="9,0")
    ap.add_argument("--glb", default="0")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--heads", type=int, default=16)
    ap.add_argument("--hd", type=int, default=80)
    a = ap.parse_args()
    lib = _lib.load()
    out = {"shape": {"batch": a.batch, "heads": a.heads, "hd": a.hd}}
    if a.win:
        out["window"] = run(lib, 14, a.batch * 25, a.heads, a.hd, [int(v) for v in a.win.split(",")], a.iters, 0)
    if a.glb:
        out["global"] = run(lib, 64, a.batch, a.heads, a.hd, [int(v) for v in a.glb.split(",")], a.iters, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
