"""Pin the CPU oracle against outputs of the unmodified reference.

The fixtures in tests/golden/*.npz were produced by tests/golden/make_golden.py,
which ran the reference's own modules (2Haff/model/segment_anything) on the
seeded weights/inputs of oracle/synth.py.  These tests run on the CPU and do
not need /root/reference.
"""
import numpy as np
import pytest
import torch

from oracle import sam_oracle as O
from oracle import synth
from tests.golden.make_golden import CASES, LITE_CASES, TAP_POS


def _run_oracle(case):
    if case in LITE_CASES:
        arch_name, first, n_frames, n_prompts, resize, original = LITE_CASES[case]
    else:
        arch_name, n_frames, n_prompts, resize, original = CASES[case]
        first = 0
    arch = synth.ARCHS[arch_name]
    w = synth.make_state_dict(arch, seed=0)
    frames = synth.make_frames(first, n_frames, valid_rows=resize[0])
    taps = {}
    emb = O.encode(w, arch, frames, O.FP32, taps=taps)
    segs = [synth.make_seg_embeds(first + i, n_prompts) for i in range(n_frames)]
    outs = O.segment(w, arch, frames, segs, [resize] * n_frames, [original] * n_frames, O.FP32, emb=emb)
    return arch, w, emb, taps, outs


def _check(case, golden, tol=2e-4):
    g = golden(case)
    arch, w, emb, taps, outs = _run_oracle(case)
    pos = torch.tensor(TAP_POS)
    for bi in range(0 if case in LITE_CASES else arch.depth):   # lite fixtures carry outputs only
        ref = torch.from_numpy(g[f"tap_block{bi}"])
        got = taps[f"block{bi}"][:, pos][:, :, pos]
        assert O.logit_rel_err(got, ref) < tol, f"block {bi}"
    assert O.logit_rel_err(emb[:, ::16], torch.from_numpy(g["emb_slice"])) < tol
    assert O.logit_rel_err(O.dense_pe(w)[0, ::32], torch.from_numpy(g["dense_pe_slice"])) < 1e-5
    for i, o in enumerate(outs):
        for side in ("left", "right"):
            ref = torch.from_numpy(g[f"f{i}_lowres_{side}"].astype(np.float32))
            lr_tol = 2e-3 if g[f"f{i}_lowres_{side}"].dtype == np.float16 else tol
            for n in range(ref.shape[0]):
                assert O.logit_rel_err(o[f"lowres_{side}"][n], ref[n]) < lr_tol
            assert np.allclose(o[f"iou_{side}"].numpy(), g[f"f{i}_iou_{side}"], atol=2e-5)
            assert np.array_equal(o[f"iou_{side}"].argmax(-1).numpy(), g[f"f{i}_iou_{side}"].argmax(-1))
            full = o[f"masks_{side}"]
            samp = torch.from_numpy(g[f"f{i}_full_{side}_samples"])
            assert O.logit_rel_err(full[:, ::37, ::41], samp) < tol
            bits = np.unpackbits(g[f"f{i}_bits_{side}"], axis=-1)[..., : full.shape[-1]].astype(bool)
            for n in range(full.shape[0]):
                assert O.mask_iou(full[n], torch.from_numpy(bits[n]).float() - 0.5) > 0.9995
        assert np.allclose(o["taxonomy"].numpy(), g[f"f{i}_taxonomy"], atol=1e-5)
        assert np.array_equal(o["taxonomy"].argmax(-1).numpy(), g[f"f{i}_taxonomy"].argmax(-1))


def test_oracle_tiny64(golden):
    _check("tiny64", golden)


def test_oracle_tiny80(golden):
    _check("tiny80", golden)


def test_oracle_vit_b_config1(golden):
    """BASELINE.json configs[0]: ViT-B, one 1024^2 frame, 2 prompts, fp32 CPU."""
    _check("vit_b_cfg1", golden)


@pytest.mark.slow
def test_oracle_vit_h_config2(golden):
    """Two frames of BASELINE.json configs[1] (ViT-H, 3 prompts)."""
    _check("vit_h_cfg2", golden)


def test_oracle_vit_l_lite(golden):
    """build_sam_vit_l, one frame x 3 prompts (outputs-only fixture)."""
    _check("vit_l_lite", golden)


def test_postprocess_matches_formula_edges():
    """Empty prompt batch and odd sizes through the two-stage bilinear resample."""
    low = torch.randn(0, 1, 256, 256)
    assert O.postprocess_masks(low, (576, 1024), (33, 77)).shape == (0, 1, 33, 77)
    low = torch.randn(2, 1, 256, 256, generator=torch.Generator().manual_seed(5))
    out = O.postprocess_masks(low, (1024, 683), (1, 1))
    assert out.shape == (2, 1, 1, 1)
    ref = torch.nn.functional.interpolate(
        torch.nn.functional.interpolate(low, (1024, 1024), mode="bilinear", align_corners=False)[..., :1024, :683],
        (1, 1), mode="bilinear", align_corners=False)
    assert torch.allclose(out, ref, atol=1e-5)


def test_train_oracle_matches_reference_autograd(golden):
    """oracle/train_oracle.py (restated forward + losses, torch.autograd on the CPU) against the unmodified
    reference's training forward / backward (tests/golden/make_golden_train.py): losses, outputs, the gradient of
    the [SEG] embeddings and the digest of every decoder parameter gradient."""
    from oracle import train_oracle as TO
    g = golden("train_tiny")
    w = synth.make_state_dict(synth.ARCHS["tiny64"], seed=0)
    emb, segs, gt_left, gt_right, tax = synth.make_train_inputs()
    case = synth.TRAIN_CASE
    r = TO.train_step(w, emb, segs, case["resize"], case["original"], gt_left, gt_right, tax)
    for k in ("loss", "mask_bce_loss", "mask_dice_loss", "taxonomy_ce_loss"):
        assert abs(float(r[k]) - float(g[k])) <= 2e-5 * max(1.0, abs(float(g[k]))), k
    for k in ("low_res_left", "low_res_right", "taxonomy"):
        assert O.logit_rel_err(r[k], torch.from_numpy(g[k])) < 2e-4, k
    for i, d in enumerate(r["d_seg"]):
        ref = torch.from_numpy(g[f"d_seg{i}"])
        assert (d - ref).abs().max() <= 2e-4 * ref.abs().max() + 1e-9, i
    names = [str(n) for n in g["param_names"]]
    assert set(names) == set(r["grads"])
    checked = 0
    for n in names:
        ref = torch.from_numpy(g["g:" + n])
        got = r["grads"][n]
        if ref.numel() == 0:          # not reached by the loss in the reference (IoU head)
            assert got is None or float(got.abs().max()) == 0.0, n
            continue
        dg = TO.grad_digest(got)
        scale = float(ref[2])         # max |grad|
        assert (dg[3:] - ref[3:]).abs().max() <= 3e-4 * scale + 1e-7, n
        # k_proj.bias gradients are mathematically zero (softmax is shift invariant): 1e-10 rounding noise both sides
        assert abs(float(dg[1] - ref[1])) <= 3e-4 * float(ref[1]) + 1e-7, n
        checked += 1
    assert checked > 100
