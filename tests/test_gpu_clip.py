"""GPU checks of the clip path (2handedafforder_b200/clip.py) with the real kernels.

One rank: what segment_clip collects (1-bit wire format, consumer thread, checksums) equals direct B200Sam.segment
calls, for a ragged clip with a variable number of [SEG] prompts per frame.  Two ranks (only when the box shows two
GPUs: `gpurun --gpus 2`): the clip is run under torchrun over NCCL and rank 0 compares every gathered frame with its
own single-GPU computation of the same frame - bit-identical masks and heads, zero checksum mismatches
(SURVEY.md 8e "Verification").  The host logic of the exchange is covered on CPU by tests/test_clip_gloo.py."""
import importlib
import json
import os
import subprocess
import sys

import pytest
import torch

from oracle import aux_oracle as A
from oracle import synth
from tests import gpu_util as G

pytestmark = pytest.mark.gpu


def test_clip_equals_direct_segment_calls():
    clip = importlib.import_module(G.PKG + ".clip")
    sam, arch, sd = G.build_sam_from_synth("tiny64", max_batch=2, max_prompts=6)
    n_frames, mb = 5, 2
    frames = synth.make_frames(0, n_frames).cuda()
    counts = [2, 0, 3, 1, 2]                       # frame 1 has no [SEG] at all (chat.py:227)
    segs = [synth.make_seg_embeds(i, 3)[:counts[i]].cuda() for i in range(n_frames)]
    resize, original = (576, 1024), (135, 243)     # 243: a row does not end on a byte boundary
    seen, batches = [], []
    host_frames = frames.cpu().pin_memory()        # loader copies from pinned host memory on the upload stream

    def load(lo, hi):
        return host_frames[lo:hi].to("cuda", non_blocking=True), segs[lo:hi]

    res = clip.segment_clip(sam.segment_bits, load, n_frames, resize, original, micro_batch=mb,
                            max_prompts_per_frame=3, prefetch=True, checksum_fn=sam.checksum64,
                            sink=lambda f, bl, br, tx, il, ir: seen.append((f, bl.shape[0], bl.is_cuda)),
                            batch_sink=lambda src, lo, cnt, bits, heads: batches.append((src, lo, list(cnt))))
    torch.cuda.synchronize()
    assert res.frames_done == n_frames and seen == [(i, counts[i], False) for i in range(n_frames)]
    assert batches == [(0, 0, [2, 0]), (0, 2, [3, 1]), (0, 4, [2])]
    assert res.checksum_mismatches == 0 and res.masks_checked == 3 * 2 * mb * 3
    for i in range(n_frames):
        ref = sam.segment(frames[i:i + 1], [segs[i]], [resize], [original], threshold=0.0)
        assert res.bits_left[i].shape == (counts[i], original[0], (original[1] + 7) // 8)
        assert torch.equal(res.masks(i, "left"), ref["masks_left"][0].cpu())
        assert torch.equal(res.masks(i, "right"), ref["masks_right"][0].cpu())
        assert torch.equal(res.taxonomy[i], ref["taxonomy"].cpu())
        assert torch.equal(res.iou_left[i], ref["iou_left"].cpu())
        assert torch.equal(res.iou_right[i], ref["iou_right"].cpu())
        # the wire format is numpy.packbits of the u8 mask, pad bits zero
        assert torch.equal(res.bits_left[i], torch.from_numpy(A.pack_bits(ref["masks_left"][0].cpu().numpy())))


def test_host_pipeline_matches_direct_segment():
    """Overlapped H2D / compute / D2H pipeline: every submitted micro-batch comes back bit-identical to a direct
    segment() call on the same inputs (4 submits over 2 slots exercise slot reuse)."""
    clip = importlib.import_module(G.PKG + ".clip")
    sam, arch, sd = G.build_sam_from_synth("tiny64", max_batch=2, max_prompts=4)
    B, NP, resize, original = 2, 2, (576, 1024), (135, 240)
    pipe = clip.HostPipeline(sam.segment_bits, B, NP, resize, original, torch.device("cuda", 0),
                             frame_dtype=torch.float32)
    batches = []
    for k in range(4):
        fr = synth.make_frames(2 * k, B).pin_memory()
        sg = torch.cat([synth.make_seg_embeds(2 * k + i, NP) for i in range(B)]).pin_memory()
        batches.append((fr, sg))
    got = []
    for k, (fr, sg) in enumerate(batches):
        pipe.submit(fr, sg)
        if k >= 1:                                  # consume the previous result while this one is in flight
            pipe.down_done[(k - 1) % 2].synchronize()
            got.append({n: v.clone() for n, v in pipe.result((k - 1) % 2).items()})
    pipe.flush()
    got.append({n: v.clone() for n, v in pipe.result(3 % 2).items()})
    assert pipe.h2d_bytes == B * 3 * 1024 * 1024 * 4 + B * NP * 256 * 4
    assert pipe.d2h_bytes == pipe.lay.nbytes < 2 * B * NP * 135 * 30 + 4096
    for k, (fr, sg) in enumerate(batches):
        segs = [sg[i * NP:(i + 1) * NP].cuda() for i in range(B)]
        ref = sam.segment(fr.cuda(), segs, [resize] * B, [original] * B, threshold=0.0)
        assert got[k]["counts"].tolist() == [NP] * B
        assert torch.equal(clip.unpack_bits(got[k]["bits"][0], 240), ref["masks_left_packed"].cpu())
        assert torch.equal(clip.unpack_bits(got[k]["bits"][1], 240), ref["masks_right_packed"].cpu())
        assert torch.equal(got[k]["heads"][0], ref["taxonomy"].cpu())
        assert torch.equal(got[k]["heads"][2], ref["iou_right"].cpu())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_two_rank_gather_is_bit_identical_to_single_gpu(tmp_path):
    """torchrun x2 over NCCL: scripts/clip_verify.py runs a ragged clip sharded over two GPUs and compares, on rank 0,
    every gathered frame with a single-GPU computation of the same frame."""
    out = tmp_path / "verify.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29617", os.path.join(G.ROOT, "scripts", "clip_verify.py"), "--arch", "tiny80",
           "--frames", "11", "--micro-batch", "2", "--out", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    rep = json.loads(out.read_text())
    assert rep["world"] == 2 and rep["frames"] == 11
    assert rep["checksum_mismatches"] == 0 and rep["masks_checked"] > 0
    assert rep["frames_bit_identical"] == 11 and rep["frames_different"] == 0
