"""GPU parity of the 1-bit wire format and the evaluation fused behind the threshold:
b200sam_postprocess_bits, b200sam_postprocess_pair, b200sam_checksum64 through the facade, against the u8 / fp32 path
of b200sam_postprocess, the numpy oracle of train_ds.py's metric functions (oracle/aux_oracle.py, pinned by
tests/golden/aux_metrics.npz) and the golden mask bits of the unmodified reference.  Byte-exact everywhere."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import aux_oracle as A
from oracle import synth
from tests import gpu_util as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tiny():
    sam, arch, sd = G.build_sam_from_synth("tiny64", max_batch=2, max_prompts=8)
    return sam


def _post_bits(sam, low, resize, original, thr=0.0):
    n, (H, W) = low.shape[0], original
    out = torch.full((n, H, (W + 7) // 8), 0xAA, dtype=torch.uint8, device="cuda")
    h = sam._ensure(low.device)
    rc = G.lib().b200sam_postprocess_bits(h, G.P(low), n, resize[0], resize[1], H, W, G.P(out), C.c_float(thr),
                                          G.stream())
    assert rc == 0, G.libmod().last_error(h)
    return out


@pytest.mark.parametrize("resize,original", [((576, 1024), (1080, 1920)), ((1024, 683), (300, 203)),
                                             ((1024, 1024), (1, 1)), ((300, 1024), (719, 2047)),
                                             ((576, 1024), (5, 4100))])
def test_bits_equal_packbits_of_the_u8_mask(tiny, resize, original):
    sam = tiny
    low = torch.randn((3, 256, 256), generator=torch.Generator().manual_seed(11)).cuda()
    for thr in (0.0, 0.4):
        logits, mask = sam._postprocess_raw(low, resize, original, True, True, thr)
        assert torch.equal(mask.bool(), logits > thr)
        bits = _post_bits(sam, low, resize, original, thr)
        torch.cuda.synchronize()
        assert np.array_equal(bits.cpu().numpy(), A.pack_bits(mask.cpu().numpy()))     # incl. zero pad bits
        assert torch.equal(sam.unpack_bits(bits, original[1]), mask)


def test_pair_bits_counts_and_gated_prediction(tiny):
    sam = tiny
    rng = np.random.default_rng(5)
    P, resize, (H, W) = 5, (576, 1024), (270, 483)
    g = torch.Generator().manual_seed(3)
    low_l = torch.randn((P, 256, 256), generator=g).cuda()
    low_r = torch.randn((P, 256, 256), generator=g).cuda() + 0.3
    tax = rng.random((P, 4)).astype(np.float32)
    tax[0] = [0.3, 0.3, 0.1, 0.1]                      # tie -> first index, like torch.argmax
    tax[1] = [0.1, 0.6, 0.2, 0.1]
    tax[2] = [0.1, 0.1, 0.2, 0.6]
    gt = (rng.random((2, H, W)) > 0.5).astype(np.uint8) * np.array([1, 255], np.uint8)[:, None, None]
    gi = [0, 1, 1, 0, 1]
    _, ml = sam._postprocess_raw(low_l, resize, (H, W), False, True, 0.1)
    _, mr = sam._postprocess_raw(low_r, resize, (H, W), False, True, 0.1)
    h = sam._ensure(low_l.device)
    Wb = (W + 7) // 8
    bl = torch.empty((P, H, Wb), dtype=torch.uint8, device="cuda")
    br = torch.empty_like(bl)
    pred = torch.empty_like(bl)
    counts = torch.full((P, 3), -7, dtype=torch.int64, device="cuda")
    taxd, gtd = torch.from_numpy(tax).cuda(), torch.from_numpy(gt).cuda()
    gid = torch.tensor(gi, dtype=torch.int32, device="cuda")
    rc = G.lib().b200sam_postprocess_pair(h, G.P(low_l), G.P(low_r), P, resize[0], resize[1], H, W, G.P(bl), G.P(br),
                                          C.c_float(0.1), G.P(taxd), G.P(gtd), G.P(gid), G.P(counts), G.P(pred),
                                          G.stream())
    assert rc == 0, G.libmod().last_error(h)
    torch.cuda.synchronize()
    assert np.array_equal(bl.cpu().numpy(), A.pack_bits(ml.cpu().numpy()))
    assert np.array_equal(br.cpu().numpy(), A.pack_bits(mr.cpu().numpy()))
    for p in range(P):
        want = A.gated_mask(ml[p].cpu().numpy(), mr[p].cpu().numpy(), tax[p])
        assert np.array_equal(pred[p].cpu().numpy(), A.pack_bits(want != 0))
        assert tuple(counts[p].tolist()) == A.mask_counts(want, gt[gi[p]])
    # the separate u8 kernel (b200sam_mask_metrics) gives the same counts
    m = sam.mask_metrics(ml, mr, taxd, gtd, gt_index=gi)
    assert torch.equal(m["counts"], counts)
    # bits only (no gt): one side may be left out; nothing to compute is an error
    bl2 = torch.empty_like(bl)
    rc = G.lib().b200sam_postprocess_pair(h, G.P(low_l), G.P(low_r), P, resize[0], resize[1], H, W, G.P(bl2), None,
                                          C.c_float(0.1), None, None, None, None, None, G.stream())
    assert rc == 0 and torch.equal(bl2, bl)
    rc = G.lib().b200sam_postprocess_pair(h, G.P(low_l), G.P(low_r), P, resize[0], resize[1], H, W, None, None,
                                          C.c_float(0.1), None, None, None, None, None, G.stream())
    assert rc == -1


def test_segment_bits_matches_segment_and_fused_metrics(tiny):
    sam = tiny
    frames = synth.make_frames(0, 2).cuda()
    segs = [synth.make_seg_embeds(0, 3).cuda(), synth.make_seg_embeds(1, 2).cuda()]
    resize, original = (576, 1024), (135, 243)
    ref = sam.segment(frames, segs, [resize] * 2, [original] * 2, threshold=0.0)
    gt = (ref["masks_left_packed"] | ref["masks_right_packed"]).contiguous()
    slot_bits = torch.zeros((2, 8, 135, 31), dtype=torch.uint8, device="cuda")
    slot_heads = torch.zeros((3, 8, 4), dtype=torch.float32, device="cuda")
    out = sam.segment_bits(frames, segs, resize, original, bits_out=slot_bits, heads_out=slot_heads, gt=gt)
    torch.cuda.synchronize()
    assert out["bits_left"].data_ptr() == slot_bits.data_ptr()            # written in place, no staging copy
    assert out["taxonomy"].data_ptr() == slot_heads.data_ptr()
    assert torch.equal(sam.unpack_bits(slot_bits[0, :5], 243), ref["masks_left_packed"])
    assert torch.equal(sam.unpack_bits(slot_bits[1, :5], 243), ref["masks_right_packed"])
    assert torch.equal(slot_heads[0, :5], ref["taxonomy"]) and torch.equal(slot_heads[1, :5], ref["iou_left"])
    assert torch.equal(slot_heads[2, :5], ref["iou_right"])
    assert not slot_bits[:, 5:].any() and not slot_heads[:, 5:].any()    # capacity beyond P untouched
    m = sam.mask_metrics(ref["masks_left_packed"], ref["masks_right_packed"], ref["taxonomy"], gt, return_pred=True)
    assert torch.equal(out["counts"], m["counts"])
    assert torch.equal(sam.unpack_bits(out["pred_bits"], 243) * 255, m["pred"])
    with pytest.raises(ValueError):
        sam.segment_bits(frames, segs, resize, original, bits_out=slot_bits[:, :4])


@pytest.mark.parametrize("shape", [(3, 21), (5, 1080, 240), (2, 8), (4, 7), (1, 100003)])
def test_checksum64_matches_oracle(tiny, shape):
    sam = tiny
    rng = np.random.default_rng(shape[-1])
    data = rng.integers(0, 256, shape, dtype=np.uint8)
    got = sam.checksum64(torch.from_numpy(data).cuda())
    assert np.array_equal(got.cpu().numpy(), A.checksum64(data))
    # position-sensitive: swapping two words changes the sum; an unaligned base pointer does not
    if data[0].size >= 16:
        sw = data.copy().reshape(shape[0], -1)
        sw[0, :8], sw[0, 8:16] = data.reshape(shape[0], -1)[0, 8:16].copy(), data.reshape(shape[0], -1)[0, :8].copy()
        if not np.array_equal(sw[0, :8], sw[0, 8:16]):
            assert A.checksum64(sw)[0] != A.checksum64(data)[0]
    flat = torch.zeros((data.size + 3,), dtype=torch.uint8, device="cuda")
    flat[3:] = torch.from_numpy(data.reshape(-1)).cuda()
    un = flat[3:].view(shape[0], -1)
    h = sam._ensure_handle_only(un.device)
    out = torch.empty((shape[0],), dtype=torch.int64, device="cuda")
    nb = un.shape[1]
    assert G.lib().b200sam_checksum64(h, G.P(un), shape[0], nb, nb, G.P(out), G.stream()) == 0
    assert np.array_equal(out.cpu().numpy(), A.checksum64(data))


def test_golden_reference_bits_in_wire_format(golden):
    """The packed masks the facade emits for ViT-B config 1 agree with the golden bits of the unmodified reference
    (stored with numpy.packbits, i.e. already in wire format) to IoU >= 0.99 per prompt."""
    from tests.golden.make_golden import CASES
    arch_name, n_frames, n_prompts, resize, original = CASES["vit_b_cfg1"]
    g = golden("vit_b_cfg1")
    sam, arch, sd = G.build_sam_from_synth(arch_name, max_batch=1, max_prompts=n_prompts)
    frames = synth.make_frames(0, 1, valid_rows=resize[0]).cuda()
    out = sam.segment_bits(frames, [synth.make_seg_embeds(0, n_prompts).cuda()], resize, original)
    for side in ("left", "right"):
        got = out[f"bits_{side}"].cpu().numpy()
        ref = g[f"f0_bits_{side}"]
        assert got.shape == ref.shape
        for n in range(n_prompts):
            a, b = np.unpackbits(got[n]), np.unpackbits(ref[n])
            iou = (a & b).sum() / max((a | b).sum(), 1)
            assert iou >= 0.99, (side, n, iou)
