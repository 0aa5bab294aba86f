#!/usr/bin/env python
"""Kernel-level breakdown of one decoder-training step (scripts/bench_train.py's workload) with torch.profiler:
prints the kernels sorted by total device time.  Measurement aid only."""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import torch.nn.functional as F
    from torch.profiler import ProfilerActivity, profile
    from oracle import synth
    dev = torch.device("cuda", 0)
    pkg = importlib.import_module("2handedafforder_b200")
    arch = synth.ARCHS["tiny64"]
    sam = pkg._build_sam(arch.dim, arch.depth, arch.heads, list(arch.global_idx), max_batch=1, max_prompts=4)
    sam.load_state_dict(synth.make_state_dict(arch, seed=0), strict=True)
    sam = sam.to(dev)
    for p in sam.parameters():
        p.requires_grad = False
    for dec in (sam.mask_decoder_left, sam.mask_decoder_right):
        dec.train()
        for p in dec.parameters():
            p.requires_grad = True
    T = sam.train_ops(dev)
    T.set_precision(args.precision)
    Fr, N, H, W = args.frames, 3, 1080, 1920
    g = torch.Generator().manual_seed(3)
    emb = (0.5 * torch.randn(Fr, 256, 64, 64, generator=g)).to(dev)
    segs = [synth.make_seg_embeds(i, N).to(dev) for i in range(Fr)]
    P = Fr * N
    gt = (torch.rand(P, H, W, generator=g) > 0.8).float().to(dev)
    w = torch.ones(P, device=dev)

    def step():
        for p in sam.parameters():
            p.grad = None
        s = [x.detach().clone().requires_grad_(True) for x in segs]
        out = sam.train_decode(emb, s)
        pl = sam.postprocess_masks(out["low_res_left"], (576, 1024), (H, W))[:, 0]
        pr = sam.postprocess_masks(out["low_res_right"], (576, 1024), (H, W))[:, 0]
        bl, dl = T.mask_losses(pl, gt, w)
        br, dr = T.mask_losses(pr, gt, w)
        loss = (bl.sum() + br.sum() + dl.sum() + dr.sum()) / P + F.cross_entropy(out["taxonomy"], out["taxonomy"].detach())
        loss.backward()
        torch.cuda.synchronize()

    step()
    step()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step()
    rows = []
    for e in prof.key_averages():
        t = getattr(e, "device_time_total", None)
        if t is None:
            t = getattr(e, "cuda_time_total", 0)
        if t > 0 and e.device_type.name == "CUDA":
            rows.append((t, e.count, e.key))
    rows.sort(reverse=True)
    tot = sum(r[0] for r in rows)
    print(f"total device time {tot / 1e3:.2f} ms over {sum(r[1] for r in rows)} kernels")
    lines = []
    for t, c, k in rows[:40]:
        line = f"{t / 1e3:9.3f} ms {100 * t / tot:5.1f}% x{c:4d}  {k[:110]}"
        print(line)
        lines.append(line)
    # strided-GEMM time by problem shape
    T.GEMM_PROFILE = []
    step()
    agg = {}
    for key, (e0, e1) in T.GEMM_PROFILE:
        a = agg.setdefault(key, [0.0, 0])
        a[0] += e0.elapsed_time(e1)
        a[1] += 1
    T.GEMM_PROFILE = None
    print("strided GEMM by shape: ms, calls, (batches, M, N, K, A(r,c), B(r,c), C(r,c))")
    for key, (ms, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:25]:
        line = f"{ms:8.3f} ms x{cnt:3d}  {key}"
        print(line)
        lines.append(line)
    if args.out:
        with open(args.out, "w") as fh:
            fh.write(f"precision {args.precision}; total device time {tot / 1e3:.2f} ms\n" + "\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
