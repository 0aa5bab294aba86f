#!/usr/bin/env python
"""Timing of the SAM branch of one training step on one GPU:

    python scripts/bench_train.py --frames 8 --prompts 3 --steps 5 --out gpurun_out/train_bench.json

Workload: F frames x N [SEG] prompts through BOTH mask decoders under autograd (B200Sam.train_decode), postprocess_masks
to 1080x1920 labels, the BCE + Dice mask losses and the taxonomy CE composed as LISA.model_forward composes them
(2Haff/model/LISA.py:284-420), backward to the gradients of every decoder parameter and of the [SEG] embeddings.
Image embeddings are synthetic (the encoder is frozen and not part of this row).

Timed with CUDA events on the current stream after warm-up.  Beside it, the same graph in plain PyTorch eager ops on
the same GPU (the torch restatement of oracle/train_oracle.py moved to CUDA: what the reference's nn.Modules would
launch; fp32 with TF32 off and on) and on the host cores for a bounded sample - measurement only, the product path never
imports the oracle.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "2handedafforder_b200"


def run(args) -> dict:
    """args: namespace with frames, prompts, steps, warmup, original, no_eager, precision, cpu_sample."""
    import torch
    import torch.nn.functional as F
    from oracle import synth
    from oracle import train_oracle as TO

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    pkg = importlib.import_module(PKG)
    arch = synth.ARCHS["tiny64"]              # the decoders are the same for every encoder size
    sam = pkg._build_sam(arch.dim, arch.depth, arch.heads, list(arch.global_idx), max_batch=1, max_prompts=4)
    sd = synth.make_state_dict(arch, seed=0)
    sam.load_state_dict(sd, strict=True)
    sam = sam.to(dev)
    for p in sam.parameters():
        p.requires_grad = False
    for dec in (sam.mask_decoder_left, sam.mask_decoder_right):
        dec.train()
        for p in dec.parameters():
            p.requires_grad = True
    T = sam.train_ops(dev)
    T.set_precision(args.precision)
    Fr, N = args.frames, args.prompts
    H, W = args.original
    resize = (576, 1024)
    g = torch.Generator().manual_seed(3)
    emb = (0.5 * torch.randn(Fr, 256, 64, 64, generator=g)).to(dev)
    segs = [synth.make_seg_embeds(i, N).to(dev) for i in range(Fr)]
    gt_l = [(torch.rand(N, H, W, generator=g) > 0.8).float().to(dev) for _ in range(Fr)]
    gt_r = [(torch.rand(N, H, W, generator=g) > 0.8).float().to(dev) for _ in range(Fr)]
    tax_gt = torch.zeros(Fr, 4)
    tax_gt[torch.arange(Fr), torch.arange(Fr) % 4] = 1.0
    tax_gt = tax_gt.to(dev)
    wl = tax_gt[:, 0] + tax_gt[:, 2] + tax_gt[:, 3]
    wr = tax_gt[:, 1] + tax_gt[:, 2] + tax_gt[:, 3]
    P = Fr * N
    gt_l_all, gt_r_all = torch.cat(gt_l), torch.cat(gt_r)
    wl_all, wr_all = wl.repeat_interleave(N), wr.repeat_interleave(N)

    def ours():
        for p in sam.parameters():
            p.grad = None
        s = [x.detach().clone().requires_grad_(True) for x in segs]
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        out = sam.train_decode(emb, s)
        ev[1].record()
        # all frames share (resize, original) here, so postprocess + losses run once over all P masks
        pl = sam.postprocess_masks(out["low_res_left"], resize, (H, W))[:, 0]
        pr = sam.postprocess_masks(out["low_res_right"], resize, (H, W))[:, 0]
        bl, dl = T.mask_losses(pl, gt_l_all, wl_all)
        br, dr = T.mask_losses(pr, gt_r_all, wr_all)
        tce = F.cross_entropy(out["taxonomy"], tax_gt.repeat_interleave(N, 0))
        loss = 2.0 * (bl.sum() + br.sum()) / (P + 1e-8) + 0.5 * (dl.sum() + dr.sum()) / (P + 1e-8) + tce
        ev[2].record()
        loss.backward()
        ev[3].record()
        torch.cuda.synchronize()
        return float(loss.detach()), [ev[i].elapsed_time(ev[i + 1]) for i in range(3)], s

    def eager(tf32: bool):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        w = {k: v.to(dev) for k, v in sd.items()}
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        r = TO.train_step(w, emb, segs, [resize] * Fr, [(H, W)] * Fr, gt_l, gt_r, tax_gt)
        t1.record()
        torch.cuda.synchronize()
        return float(r["loss"]), t0.elapsed_time(t1)

    n0 = sam.launch_count()
    for _ in range(args.warmup):
        loss, _, _ = ours()
    per_step_launches = (sam.launch_count() - n0) // max(args.warmup, 1)
    runs = [ours() for _ in range(args.steps)]
    phases = [sorted(r[1][i] for r in runs)[len(runs) // 2] for i in range(3)]
    wall0 = time.perf_counter()
    ours()
    wall = (time.perf_counter() - wall0) * 1e3
    rep = {"workload": f"{Fr} frames x {N} prompts, both decoders, postprocess to {H}x{W}, BCE+Dice+taxonomy CE, backward",
           "prompts_total": P, "precision": args.precision, "loss": runs[-1][0],
           "ms": {"decoders_forward": phases[0], "postprocess_and_losses": phases[1], "backward": phases[2],
                  "step_device": sum(phases), "step_wall": wall},
           "gpu_launches_per_step": int(per_step_launches), "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
    if not args.no_eager:
        torch.cuda.reset_peak_memory_stats()
        for tf32 in (False, True):
            eager(tf32)
            vals = sorted(eager(tf32) for _ in range(3))
            rep[f"torch_eager_{'tf32' if tf32 else 'fp32'}_ms"] = vals[1][1]
            rep[f"torch_eager_{'tf32' if tf32 else 'fp32'}_loss"] = vals[1][0]
        rep["torch_eager_peak_mem_gb"] = torch.cuda.max_memory_allocated() / 2 ** 30
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False
    if args.cpu_sample > 0:
        k = args.cpu_sample
        torch.set_num_threads(os.cpu_count())
        cpu = lambda xs: [x.cpu() for x in xs]
        t0 = time.perf_counter()
        r = TO.train_step(sd, emb[:k].cpu(), cpu(segs[:k]), [resize] * k, [(H, W)] * k, cpu(gt_l[:k]), cpu(gt_r[:k]),
                          tax_gt[:k].cpu())
        dt = time.perf_counter() - t0
        rep["cpu_oracle"] = {"frames": k, "prompts": k * N, "seconds": dt, "cores": os.cpu_count(),
                             "ms_per_prompt": dt * 1e3 / (k * N)}
    rep["ms_per_prompt"] = rep["ms"]["step_device"] / P
    T.set_precision("tf32")
    for dec in (sam.mask_decoder_left, sam.mask_decoder_right):
        dec.eval()
    return rep


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--prompts", type=int, default=3)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--original", type=int, nargs=2, default=[1080, 1920])
    ap.add_argument("--no-eager", action="store_true")
    ap.add_argument("--precision", default="tf32", choices=["tf32", "fp32"],
                    help="arithmetic of the big Linear layers (train.set_precision)")
    ap.add_argument("--cpu-sample", type=int, default=1, help="frames of the CPU oracle sample (0 = skip)")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    rep = run(args)
    print(json.dumps(rep))
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as fh:
            json.dump(rep, fh)
    return 0


if __name__ == "__main__":
    sys.exit(main())
