"""End-to-end parity of the CUDA path (through the facade -> C ABI) against
(a) the committed golden outputs of the UNMODIFIED reference (tests/golden/*.npz, fp32 CPU) and
(b) the CPU oracle on the same seeded inputs.

Pass criteria (BASELINE.json north_star): per mask max|dlogit| / max|ref logit| <= 2e-2, thresholded-mask
IoU >= 0.99 per prompt, exact argmax agreement of the IoU head and of the taxonomy head.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import sam_oracle as O
from oracle import synth
from tests import gpu_util as G
from tests.golden.make_golden import CASES, TAP_POS

pytestmark = pytest.mark.gpu

REPORT = os.path.join(G.ROOT, "gpurun_out", "parity_report.jsonl")


def _report(**kw):
    try:
        os.makedirs(os.path.dirname(REPORT), exist_ok=True)
        with open(REPORT, "a") as fh:
            fh.write(json.dumps(kw) + "\n")
    except OSError:
        pass


def _taps(sam, frames, nblocks):
    lib, h = G.lib(), sam._handle
    assert lib.b200sam_debug_stop_after_block(h, nblocks) == 0
    try:
        sam._encode(frames, want_output=False)
        B, D = frames.shape[0], sam.arch[0]
        out = torch.empty((B, 64, 64, D), dtype=torch.float32, device="cuda")
        assert lib.b200sam_debug_read_tokens(h, G.P(out), B, G.stream()) == 0
        torch.cuda.synchronize()
    finally:
        lib.b200sam_debug_stop_after_block(h, -1)
    return out


def _run_case(case, golden, tap_blocks, frame_dtype=torch.float32):
    arch_name, n_frames, n_prompts, resize, original = CASES[case]
    g = golden(case)
    sam, arch, sd = G.build_sam_from_synth(arch_name, max_batch=n_frames, max_prompts=n_frames * n_prompts)
    frames = synth.make_frames(0, n_frames, valid_rows=resize[0]).cuda().to(frame_dtype)
    segs = [synth.make_seg_embeds(i, n_prompts).cuda() for i in range(n_frames)]
    sam._ensure(frames.device, n_frames, n_frames * n_prompts)
    pos = torch.tensor(TAP_POS).cuda()
    worst = {}
    for nb in tap_blocks:
        got = _taps(sam, frames, nb)[:, pos][:, :, pos].cpu()
        ref = torch.from_numpy(g[f"tap_block{nb - 1}"])
        err = O.logit_rel_err(got, ref)
        worst[f"block{nb - 1}"] = err
        _report(case=case, stage=f"block{nb - 1}", rel_err=err)
        assert err < 3e-2, f"{case}: residual stream after block {nb - 1} off by {err:.3e}"
    emb = sam.image_encoder(frames)
    err = O.logit_rel_err(emb[:, ::16].float().cpu(), torch.from_numpy(g["emb_slice"]))
    _report(case=case, stage="emb", rel_err=err)
    assert err < 3e-2, f"{case}: image embedding off by {err:.3e}"
    pe = sam.prompt_encoder.get_dense_pe()
    assert O.logit_rel_err(pe[0, ::32].cpu(), torch.from_numpy(g["dense_pe_slice"])) < 1e-4

    out = sam.segment(frames, segs, [resize] * n_frames, [original] * n_frames, threshold=0.0, return_logits=True)
    torch.cuda.synchronize()
    offs = out["offsets"]
    for i in range(n_frames):
        sl = slice(offs[i], offs[i + 1])
        for side in ("left", "right"):
            ref_low = torch.from_numpy(g[f"f{i}_lowres_{side}"].astype(np.float32))
            got_low = out[f"lowres_{side}"][sl].cpu()
            logits = out[f"logits_{side}"][i].cpu()
            masks = out[f"masks_{side}"][i].cpu()
            assert torch.equal(masks.bool(), logits > 0)
            bits = np.unpackbits(g[f"f{i}_bits_{side}"], axis=-1)[..., : logits.shape[-1]].astype(bool)
            samp = torch.from_numpy(g[f"f{i}_full_{side}_samples"])
            absmax = torch.from_numpy(g[f"f{i}_full_{side}_absmax"])
            for n in range(n_prompts):
                e_low = O.logit_rel_err(got_low[n], ref_low[n])
                e_full = float((logits[n, ::37, ::41] - samp[n]).abs().max() / absmax[n])
                iou = O.mask_iou(logits[n], torch.from_numpy(bits[n]).float() - 0.5)
                _report(case=case, frame=i, side=side, prompt=n, lowres_err=e_low, full_err=e_full, iou=iou)
                assert e_low <= 2e-2 + 2e-3, f"{case} f{i} {side} p{n}: low-res logit error {e_low:.3e}"
                assert e_full <= 2e-2, f"{case} f{i} {side} p{n}: full-res logit error {e_full:.3e}"
                assert iou >= 0.99, f"{case} f{i} {side} p{n}: thresholded IoU {iou:.4f}"
            iou_head = out[f"iou_{side}"][sl].cpu().numpy()
            _report(case=case, frame=i, side=side, iou_head_err=float(np.abs(iou_head - g[f"f{i}_iou_{side}"]).max()))
            assert np.array_equal(iou_head.argmax(-1), g[f"f{i}_iou_{side}"].argmax(-1))
        tax = out["taxonomy"][sl].cpu().numpy()
        _report(case=case, frame=i, tax_err=float(np.abs(tax - g[f"f{i}_taxonomy"]).max()))
        assert np.array_equal(tax.argmax(-1), g[f"f{i}_taxonomy"].argmax(-1))
        assert np.allclose(tax.sum(-1), 1.0, atol=1e-5)
    return sam, out


def test_path_tiny64(golden):
    _run_case("tiny64", golden, tap_blocks=[1, 2])


def test_path_tiny80(golden):
    _run_case("tiny80", golden, tap_blocks=[1, 2, 3])


def test_path_vit_b_config1(golden):
    """BASELINE.json configs[0]: ViT-B, one frame, 2 prompts."""
    _run_case("vit_b_cfg1", golden, tap_blocks=[1, 3, 6, 12])


def test_path_vit_h_config2(golden):
    """Two frames of BASELINE.json configs[1] (ViT-H, 3 prompts per frame), bf16 frames as inference.py feeds them."""
    _run_case("vit_h_cfg2", golden, tap_blocks=[1, 8, 16, 32], frame_dtype=torch.bfloat16)
