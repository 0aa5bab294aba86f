#!/usr/bin/env python
"""On-hardware proof that the N-GPU gathered clip equals the single-GPU result (SURVEY.md 8e "Verification",
BASELINE configs 3 and 5).  Launch under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29617 \
        scripts/clip_verify.py --arch vit_h --frames 2048 --micro-batch 8 --out gpurun_out/clip_verify_n8.json

Every rank segments its contiguous shard of a synthetic clip (frame i is a pure function of i) and the packed masks
+ heads are gathered to rank 0 through clip.segment_clip, streamed to the host (keep=False) into a consumer that keeps
only a 64-bit digest per frame.  Rank 0 then recomputes EVERY frame of the clip by itself, on its own GPU in plain
micro-batches, and compares digests: `frames_bit_identical` must equal `frames`.  The transport checksums
(b200sam_checksum64, producer side vs. what the gather delivered) are reported as well.
"""
from __future__ import annotations

import argparse
import hashlib
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "2handedafforder_b200"
# build_sam.py:15-48 plus the two small test shapes: (dim, depth, heads, global attention blocks)
ARCHS = {"tiny64": (256, 2, 4, (1,)), "tiny80": (320, 3, 4, (2,)), "vit_b": (768, 12, 12, (2, 5, 8, 11)),
         "vit_l": (1024, 24, 16, (5, 11, 17, 23)), "vit_h": (1280, 32, 16, (7, 15, 23, 31))}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="tiny80", choices=["tiny64", "tiny80", "vit_b", "vit_l", "vit_h"])
    ap.add_argument("--frames", type=int, default=11)
    ap.add_argument("--micro-batch", type=int, default=2)
    ap.add_argument("--prompts", type=int, default=3, help="max [SEG] prompts per frame")
    ap.add_argument("--ragged", type=int, default=1, help="1: frame i has (i*7+1)%%(prompts+1) prompts; 0: always max")
    ap.add_argument("--original", type=int, nargs=2, default=[1080, 1920])
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pkg = importlib.import_module(PKG)
    clip = importlib.import_module(PKG + ".clip")
    dim, depth, heads, gidx = ARCHS[args.arch]
    mb, NP = args.micro_batch, args.prompts
    # the facade's own seeded init (identical on every rank); pos_embed / rel_pos_* are zeros at init in the
    # reference (image_encoder.py:70-74,232-233) and get N(0, 0.02) noise so that the rel-pos path does real arithmetic
    sam = pkg._build_sam(dim, depth, heads, list(gidx), max_batch=mb, max_prompts=mb * NP)
    g0 = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for name, p in sam.named_parameters():
            if "pos_embed" in name or "rel_pos" in name:
                p.copy_(0.02 * torch.randn(p.shape, generator=g0))
    sam = sam.to(dev)
    sam.freeze_weights(True)
    resize, original = (576, 1024), tuple(args.original)
    dt = torch.bfloat16 if args.arch in ("vit_h", "vit_l") else torch.float32

    def n_prompts(i: int) -> int:
        return (i * 7 + 1) % (NP + 1) if args.ragged else NP

    def frame(i: int):
        g = torch.Generator(device=dev).manual_seed(10 + i)
        f = torch.randn((3, 1024, 1024), generator=g, device=dev)
        f[:, resize[0]:, :] = 0.0
        return f.to(dt)

    def seg(i: int):
        g = torch.Generator(device=dev).manual_seed(100003 + i)
        return torch.randn((NP, 256), generator=g, device=dev)[: n_prompts(i)]

    def load(lo, hi):
        return torch.stack([frame(i) for i in range(lo, hi)]), [seg(i) for i in range(lo, hi)]

    def digest(bl, br, tx, il, ir) -> str:
        h = hashlib.blake2b(digest_size=8)
        for t in (bl, br, tx, il, ir):
            h.update(t.contiguous().numpy().tobytes())
        return h.hexdigest()

    got = {}
    stats = {"batches": 0, "bytes": 0}

    def sink(f, bl, br, tx, il, ir):
        got[f] = (int(bl.shape[0]), digest(bl, br, tx, il, ir))

    def batch_sink(src, lo, cnt, bits, heads):
        stats["batches"] += 1
        stats["bytes"] += bits.numel() + heads.numel() * 4

    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = clip.segment_clip(sam.segment_bits, load, args.frames, resize, original, micro_batch=mb,
                            max_prompts_per_frame=NP, device=dev, sink=sink, batch_sink=batch_sink, keep=False,
                            checksum_fn=sam.checksum64)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t0

    rc = 0
    if rank == 0:
        same = diff = 0
        bad = []
        for lo in range(0, args.frames, mb):
            hi = min(lo + mb, args.frames)
            fr, sg = load(lo, hi)
            out = sam.segment_bits(fr, sg, resize, original, threshold=0.0)
            offs = out["offsets"]
            host = {k: out[k].cpu() for k in ("bits_left", "bits_right", "taxonomy", "iou_left", "iou_right")}
            for j, f in enumerate(range(lo, hi)):
                sl = slice(offs[j], offs[j + 1])
                d = digest(*(host[k][sl] for k in ("bits_left", "bits_right", "taxonomy", "iou_left", "iou_right")))
                ok = f in got and got[f] == (offs[j + 1] - offs[j], d)
                same += ok
                diff += not ok
                if not ok and len(bad) < 8:
                    bad.append(f)
        rep = {"world": world, "arch": args.arch, "frames": args.frames, "micro_batch": mb,
               "prompts_per_frame": "ragged 0..%d" % NP if args.ragged else NP, "original": list(original),
               "frames_gathered": res.frames_done, "frames_bit_identical": same, "frames_different": diff,
               "first_bad_frames": bad, "masks_checked": res.masks_checked,
               "checksum_mismatches": res.checksum_mismatches, "gather_batches": stats["batches"],
               "gathered_bytes": stats["bytes"], "clip_wall_s": wall, "clip_frames_per_s": args.frames / wall,
               "note": "frames generated on the device inside the timed region; wall clock incl. generation"}
        print(json.dumps(rep))
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as fh:
                json.dump(rep, fh)
        rc = 0 if (diff == 0 and res.checksum_mismatches == 0 and res.frames_done == args.frames) else 1
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
