#!/usr/bin/env python
"""Experiment: does decode + postprocess of batch k overlap with the encode of batch k+1?
ViT-H, 8 frames x 3 prompts, device-resident inputs.  Serial = segment_bits back to back on one stream.
Overlapped = encode on the main stream (embedding returned as an NCHW tensor), decode + postprocess on a second
stream behind an event; the decoder then reads its own token-major copy (emb_in), so the next encode may overwrite
emb_tok.  CUDA events over 10 steps after 3 warm-ups."""
import ctypes as C
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

pkg = importlib.import_module("2handedafforder_b200")
_lib = importlib.import_module("2handedafforder_b200._lib")
dev = torch.device("cuda", 0)
B, NP, H, W = 8, 3, 1080, 1920
sam = pkg.build_sam_vit_h(None) if hasattr(pkg, "build_sam_vit_h") else None
if sam is None:
    sam = pkg._build_sam(1280, 32, 16, [7, 15, 23, 31], max_batch=B, max_prompts=B * NP)
g0 = torch.Generator().manual_seed(1)
with torch.no_grad():
    for name, p in sam.named_parameters():
        if "pos_embed" in name or "rel_pos" in name:
            p.copy_(0.02 * torch.randn(p.shape, generator=g0))
sam = sam.to(dev)
sam.freeze_weights(True)
frames = torch.randn((B, 3, 1024, 1024), device=dev).to(torch.bfloat16)
segs = [torch.randn((NP, 256), device=dev) for _ in range(B)]
resize, orig = (576, 1024), (H, W)
Wb = (W + 7) // 8
P = B * NP
offsets = [i * NP for i in range(B + 1)]
seg_cat = torch.cat(segs, 0)


def serial(n):
    for _ in range(n):
        sam.segment_bits(frames, segs, resize, orig)


main = torch.cuda.current_stream()
side = torch.cuda.Stream()
bits = [torch.empty((2, P, H, Wb), dtype=torch.uint8, device=dev) for _ in range(2)]


def overlapped(n):
    evs = []
    for k in range(n):
        emb = sam.get_visual_embs(frames)                     # main stream
        ev = torch.cuda.Event()
        ev.record(main)
        with torch.cuda.stream(side):
            side.wait_event(ev)
            res = sam._decode_raw(emb, B, seg_cat, offsets, True, True, dev)
            h = sam._handle
            rc = _lib.load().b200sam_postprocess_pair(
                h, C.c_void_p(res["lowres_left"].data_ptr()), C.c_void_p(res["lowres_right"].data_ptr()), P, resize[0],
                resize[1], H, W, C.c_void_p(bits[k & 1][0].data_ptr()), C.c_void_p(bits[k & 1][1].data_ptr()),
                C.c_float(0.0), None, None, None, None, None, C.c_void_p(side.cuda_stream))
            assert rc == 0
            emb.record_stream(side)
            for t in res.values():
                t.record_stream(side)
    main.wait_stream(side)


def timed(fn, n=10):
    fn(3)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn(n)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.no_grad():
    rep = {}
    for rnd in range(2):
        rep[f"serial_ms_{rnd}"] = timed(serial)
        rep[f"overlapped_ms_{rnd}"] = timed(overlapped)
    # encode only, for scale
    rep["encode_only_ms"] = timed(lambda n: [sam.get_visual_embs(frames) for _ in range(n)])
    # the overlapped result equals the serial one
    ref = sam.segment_bits(frames, segs, resize, orig)
    overlapped(2)
    torch.cuda.synchronize()
    rep["bit_identical"] = bool(torch.equal(bits[1][0], ref["bits_left"]) and torch.equal(bits[1][1], ref["bits_right"]))
print(json.dumps(rep, indent=1))
if len(sys.argv) > 1:
    json.dump(rep, open(sys.argv[1], "w"), indent=1)
