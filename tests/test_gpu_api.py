"""The reference-facing call surface on the GPU: the exact call sequence of LISA.evaluate
(2Haff/model/LISA.py:487-532) against the fused `segment()` fast path, edge cases, postprocess parity."""
import pytest
import torch
import torch.nn.functional as F

from oracle import sam_oracle as O
from oracle import synth
from tests import gpu_util as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tiny():
    sam, arch, sd = G.build_sam_from_synth("tiny64", max_batch=2, max_prompts=8)
    return sam, arch, sd


def _lisa_sequence(sam, frames, segs, resize, original):
    """Verbatim shape of the reference loop: B=1 encodes, per-image prompt encoder + both decoders + postprocess."""
    embs = torch.cat([sam.image_encoder(frames[i:i + 1]) for i in range(frames.shape[0])], 0)   # LISA.py:160-166
    res = []
    for i in range(frames.shape[0]):
        sparse, dense = sam.prompt_encoder(points=None, boxes=None, masks=None, text_embeds=segs[i].unsqueeze(1))
        sparse = sparse.to(segs[i].dtype)
        low_l, iou_l, tax = sam.mask_decoder_left(
            image_embeddings=embs[i].unsqueeze(0), image_pe=sam.prompt_encoder.get_dense_pe(),
            sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense, multimask_output=False)
        pl = sam.postprocess_masks(low_l, input_size=resize, original_size=original)
        low_r, iou_r = sam.mask_decoder_right(
            image_embeddings=embs[i].unsqueeze(0), image_pe=sam.prompt_encoder.get_dense_pe(),
            sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense, multimask_output=False)
        pr = sam.postprocess_masks(low_r, input_size=resize, original_size=torch.Size(original))   # LISA.py:309
        res.append((pl[:, 0], pr[:, 0], tax, iou_l, iou_r))
    return res


def test_lisa_call_sequence_matches_fused_path(tiny):
    sam, arch, _ = tiny
    frames = synth.make_frames(0, 2).cuda()
    segs = [synth.make_seg_embeds(0, 3).cuda(), synth.make_seg_embeds(1, 1).cuda()]   # ragged prompt counts
    resize, original = (576, 1024), (270, 480)
    seq = _lisa_sequence(sam, frames, segs, resize, original)
    out = sam.segment(frames, segs, [resize] * 2, [original] * 2, return_logits=True)
    offs = out["offsets"]
    assert offs == [0, 3, 4]
    for i, (pl, pr, tax, iou_l, iou_r) in enumerate(seq):
        assert pl.dtype == torch.float32 and pl.shape == (segs[i].shape[0], 270, 480)
        # B=1 encodes vs one batched encode: reduction order must not depend on the batch (Appendix A13)
        assert torch.equal(pl, out["logits_left"][i])
        assert torch.equal(pr, out["logits_right"][i])
        sl = slice(offs[i], offs[i + 1])
        assert torch.equal(tax, out["taxonomy"][sl])
        assert torch.equal(iou_l, out["iou_left"][sl][:, 0:1])
        assert torch.equal(iou_r, out["iou_right"][sl][:, 0:1])


def test_zero_prompts(tiny):
    """No [SEG] generated -> empty tensors, not an error (chat.py:227, Appendix A14)."""
    sam, _, _ = tiny
    frames = synth.make_frames(0, 1).cuda()
    emb = sam.image_encoder(frames)
    sparse, dense = sam.prompt_encoder(None, None, None, torch.empty((0, 1, 256), device="cuda"))
    assert sparse.shape == (0, 1, 256) and dense.shape == (0, 256, 64, 64)
    low, iou, tax = sam.mask_decoder_left(emb, sam.prompt_encoder.get_dense_pe(), sparse, dense, False)
    assert low.shape == (0, 1, 256, 256) and iou.shape == (0, 1) and tax.shape == (0, 4)
    assert sam.postprocess_masks(low, (576, 1024), (33, 77)).shape == (0, 1, 33, 77)
    out = sam.segment(frames, [torch.empty((0, 256), device="cuda")], [(576, 1024)], [(33, 77)])
    assert out["masks_left"][0].shape == (0, 33, 77)


def test_unsupported_inputs_raise(tiny):
    sam, _, _ = tiny
    with pytest.raises(NotImplementedError):
        sam.prompt_encoder(points=(torch.zeros(1, 1, 2), torch.zeros(1, 1)), boxes=None, masks=None, text_embeds=None)
    with pytest.raises(ValueError):
        sam.image_encoder(torch.zeros((1, 3, 512, 512), device="cuda"))
    with pytest.raises(RuntimeError):
        sam.image_encoder.forward(torch.zeros((1, 3, 1024, 1024)))   # CPU tensor: no fallback
    emb = torch.zeros((1, 256, 64, 64), device="cuda")
    with pytest.raises(NotImplementedError):
        sam.mask_decoder_left(emb, None, torch.zeros((1, 1, 256), device="cuda"), None, multimask_output=True)


def test_weights_follow_parameter_updates(tiny):
    """`.bfloat16()` / in-place edits of the nn.Parameters are picked up (lazy repack)."""
    sam, _, _ = tiny
    frames = synth.make_frames(3, 1).cuda()
    a = sam.image_encoder(frames).clone()
    p = sam.image_encoder.neck._modules["3"].bias
    keep = p.detach().clone()
    with torch.no_grad():
        p.add_(1.0)
    b = sam.image_encoder(frames)
    with torch.no_grad():
        p.copy_(keep)
    assert torch.allclose(b, a + 1.0, atol=1e-5)
    assert torch.equal(sam.image_encoder(frames), a)
    # a LayerNorm weight of the encoder: the folded qkv / lin1 weights (gamma . W, c, d) are derived buffers and must be
    # rebuilt; the stand-alone LayerNorm path (debug variant 5 = 1) is the reference for the new value
    blk = sam.image_encoder.blocks._modules["0"]
    keep_g, keep_b = blk.norm1.weight.detach().clone(), blk.norm2.bias.detach().clone()
    try:
        with torch.no_grad():
            blk.norm1.weight.mul_(1.5)
            blk.norm2.bias.add_(0.25)
        folded = sam.image_encoder(frames).clone()
        assert G.lib().b200sam_debug_set_variant(5, 1) == 0
        standalone = sam.image_encoder(frames).clone()
        assert G.lib().b200sam_debug_set_variant(5, 0) == 0
        assert G.rel_err(folded, a) > 1e-2                       # the update is visible
        assert G.rel_err(folded, standalone) < 1.5e-2            # and both LayerNorm paths see the same new weights
    finally:
        G.lib().b200sam_debug_set_variant(5, 0)
        with torch.no_grad():
            blk.norm1.weight.copy_(keep_g)
            blk.norm2.bias.copy_(keep_b)
    assert torch.equal(sam.image_encoder(frames), a)


@pytest.mark.parametrize("resize,original", [((576, 1024), (1080, 1920)), ((1024, 683), (300, 200)),
                                             ((1024, 1024), (1, 1)), ((300, 1024), (719, 2047))])
def test_postprocess_matches_interpolate(tiny, resize, original):
    sam, _, _ = tiny
    low = torch.randn((3, 1, 256, 256), generator=torch.Generator().manual_seed(5)).cuda()
    got = sam.postprocess_masks(low, resize, original)
    ref = F.interpolate(low, (1024, 1024), mode="bilinear", align_corners=False)[..., : resize[0], : resize[1]]
    ref = F.interpolate(ref, original, mode="bilinear", align_corners=False)
    assert got.shape == ref.shape and got.dtype == torch.float32
    assert float((got - ref).abs().max()) < 2e-5
    cpu = O.postprocess_masks(low.cpu(), resize, original)
    assert float((got.cpu() - cpu).abs().max()) < 1e-4   # fp32 FMA-contraction differences only (|logit| ~ 4)
    _, mask = sam._postprocess_raw(low[:, 0].contiguous(), resize, original, False, True, 0.25)
    logits, _ = sam._postprocess_raw(low[:, 0].contiguous(), resize, original, True, False, 0.0)
    assert torch.equal(mask.bool(), logits > 0.25)


def test_training_forward_gets_a_graph_and_frozen_parts_refuse(tiny):
    """LISA.model_forward with train_mask_decoder=True (LISA.py:84-90,284-325): the decoders and postprocess_masks
    return tensors WITH grad_fn (the fp32 training kernels, tests/test_gpu_train.py checks the numbers); the frozen
    encoder and the fused inference fast path refuse grad-requiring inputs instead of returning detached tensors;
    inference calls are unaffected."""
    sam, _, _ = tiny
    frames = synth.make_frames(0, 1).cuda()
    with torch.no_grad():
        emb = sam.image_encoder(frames)
    seg = synth.make_seg_embeds(0, 2).cuda().requires_grad_(True)          # output of the trainable text_hidden_fcs
    sparse, dense = sam.prompt_encoder(None, None, None, seg.unsqueeze(1))
    assert sparse.requires_grad
    low_t, iou_t, tax_t = sam.mask_decoder_left(emb, sam.prompt_encoder.get_dense_pe(), sparse, dense, False)
    assert low_t.grad_fn is not None and tax_t.grad_fn is not None and low_t.shape == (2, 1, 256, 256)
    with torch.no_grad():                                                   # LISA.evaluate / validate
        low, iou, tax = sam.mask_decoder_left(emb, sam.prompt_encoder.get_dense_pe(), sparse, dense, False)
    assert low.shape == (2, 1, 256, 256) and low.grad_fn is None
    # same weights, fp32 training kernels vs the TF32 inference kernels
    assert G.rel_err(low_t.detach(), low) < 2e-2
    up = sam.postprocess_masks(low.clone().requires_grad_(True), (576, 1024), (64, 64))
    assert up.grad_fn is not None
    with pytest.raises(NotImplementedError, match="no backward"):
        sam.image_encoder(frames.clone().requires_grad_(True))
    out = sam.segment(frames, [seg], [(576, 1024)], [(64, 64)])            # the fused fast path is @torch.no_grad()
    assert out["taxonomy"].grad_fn is None
    sam.mask_decoder_right.train()                                          # LISA.py:86-90
    try:
        for p in sam.mask_decoder_right.parameters():
            p.requires_grad = True
        low_r, _ = sam.mask_decoder_right(emb, sam.prompt_encoder.get_dense_pe(), sparse.detach(), dense, False)
        assert low_r.grad_fn is not None
    finally:
        sam.mask_decoder_right.eval()
    low_e, _ = sam.mask_decoder_right(emb, sam.prompt_encoder.get_dense_pe(), sparse.detach(), dense, False)   # eval
    assert low_e.grad_fn is None


def test_weight_resync_keeps_memory_flat(tiny):
    """validate() between training epochs of the mask decoders (train_ds.py:236-244,625-758): every parameter change
    re-finalises the handle; the derived buffers must be replaced, not accumulated, and the outputs must follow."""
    sam, _, _ = tiny
    frames = synth.make_frames(1, 1).cuda()
    seg = [synth.make_seg_embeds(1, 2).cuda()]
    args = ([(576, 1024)], [(64, 96)])
    p = sam.mask_decoder_left.iou_prediction_head.layers._modules["2"].bias
    q = sam.image_encoder.blocks._modules["1"].attn.rel_pos_h
    base = sam.segment(frames, seg, *args)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for it in range(20):
        with torch.no_grad():
            p.add_(0.5)
            if it % 5 == 0:
                q.mul_(1.0)                     # encoder group: version bump -> encoder packs rebuilt too
        out = sam.segment(frames, seg, *args)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert abs(free0 - free1) < 64 << 20, f"free memory moved by {(free0 - free1) / 2**20:.1f} MiB over 20 re-syncs"
    assert torch.allclose(out["iou_left"], base["iou_left"] + 10.0, atol=1e-4)
    assert torch.equal(out["masks_left_packed"], base["masks_left_packed"])
    with torch.no_grad():
        p.sub_(10.0)


def test_fp16_overflow_of_relpos_projection_is_reported(tiny):
    """The global blocks keep q.Rh / q.Rw as fp16: values beyond 65504 are saturated (no inf/NaN in the output) and the
    NEXT library call fails loudly; restoring the weights restores service."""
    sam, _, _ = tiny
    frames = synth.make_frames(2, 1).cuda()
    good = sam.image_encoder(frames).clone()
    rel = sam.image_encoder.blocks._modules["1"].attn.rel_pos_w          # tiny64: block 1 is the global one
    keep = rel.detach().clone()
    with torch.no_grad():
        rel.fill_(3.0e4)
    bad = sam.image_encoder(frames)
    torch.cuda.synchronize()
    assert torch.isfinite(bad.float()).all()
    with pytest.raises(RuntimeError, match="saturated the fp16 rel-pos"):
        sam.image_encoder(frames)
    with torch.no_grad():
        rel.copy_(keep)
    assert torch.equal(sam.image_encoder(frames), good)
