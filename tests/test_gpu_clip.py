"""GPU check of the clip path (2handedafforder_b200/clip.py) with the real kernels on one rank: what segment_clip
collects equals direct B200Sam.segment calls, for a ragged clip with a variable number of [SEG] prompts per frame
(the multi-rank exchange itself is covered by tests/test_clip_gloo.py on CPU and by `bench.py --gpus N`)."""
import importlib

import pytest
import torch

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
    resize, original = (576, 1024), (135, 240)
    seen = []
    host_frames = frames.cpu().pin_memory()        # loader copies from pinned host memory on the upload stream

    def load(lo, hi):
        return host_frames[lo:hi].to("cuda", non_blocking=True), segs[lo:hi]

    res = clip.segment_clip(sam.segment, load, n_frames, resize, original,
                            micro_batch=mb, max_prompts_per_frame=3, prefetch=True,
                            sink=lambda f, ml, mr, tx, il, ir: seen.append((f, ml.shape[0])))
    torch.cuda.synchronize()
    assert res.frames_done == n_frames and seen == [(i, counts[i]) for i in range(n_frames)]
    for i in range(n_frames):
        ref = sam.segment(frames[i:i + 1], [segs[i]], [resize], [original], threshold=0.0)
        assert res.masks_left[i].shape == (counts[i], *original)
        assert torch.equal(res.masks_left[i], ref["masks_left"][0])
        assert torch.equal(res.masks_right[i], ref["masks_right"][0])
        assert torch.equal(res.taxonomy[i], ref["taxonomy"])
        assert torch.equal(res.iou_left[i], ref["iou_left"]) and torch.equal(res.iou_right[i], ref["iou_right"])


def test_host_pipeline_matches_direct_segment():
    """Overlapped H2D / compute / D2H pipeline: every submitted micro-batch comes back bit-identical to a direct
    segment() call on the same inputs (4 submits over 2 slots exercise slot reuse)."""
    clip = importlib.import_module(G.PKG + ".clip")
    sam, arch, sd = G.build_sam_from_synth("tiny64", max_batch=2, max_prompts=4)
    B, NP, resize, original = 2, 2, (576, 1024), (135, 240)
    pipe = clip.HostPipeline(sam.segment, B, NP, resize, original, torch.device("cuda", 0), frame_dtype=torch.float32)
    batches = []
    for k in range(4):
        fr = synth.make_frames(2 * k, B).pin_memory()
        sg = torch.cat([synth.make_seg_embeds(2 * k + i, NP) for i in range(B)]).pin_memory()
        batches.append((fr, sg))
    got = []
    for k, (fr, sg) in enumerate(batches):
        s = pipe.submit(fr, sg)
        if k >= 1:                                  # consume the previous result while this one is in flight
            pipe.down_done[(k - 1) % 2].synchronize()
            got.append({n: v.clone() for n, v in pipe.host_out[(k - 1) % 2].items()})
    pipe.flush()
    got.append({n: v.clone() for n, v in pipe.host_out[3 % 2].items()})
    assert pipe.h2d_bytes == B * 3 * 1024 * 1024 * 4 + B * NP * 256 * 4
    for k, (fr, sg) in enumerate(batches):
        segs = [sg[i * NP:(i + 1) * NP].cuda() for i in range(B)]
        ref = sam.segment(fr.cuda(), segs, [resize] * B, [original] * B, threshold=0.0)
        assert torch.equal(got[k]["masks"][0], ref["masks_left_packed"].cpu())
        assert torch.equal(got[k]["masks"][1], ref["masks_right_packed"].cpu())
        assert torch.equal(got[k]["heads"][0], ref["taxonomy"].cpu())
