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
