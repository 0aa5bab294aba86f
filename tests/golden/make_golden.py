"""Generate the committed golden fixtures by running the UNMODIFIED reference.

Run in the build container only (the reference lives at /root/reference, which
does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py [case ...]

For each case the seeded weights of oracle/synth.py are loaded with
`load_state_dict(strict=True)` into the reference's own modules
(2Haff/model/segment_anything: build_sam_vit_b/h, or `_build_sam` for the tiny
test shapes), the reference's call sequence of 2Haff/model/LISA.py:487-532 is
executed in fp32 on the CPU, and the outputs are written to
tests/golden/<case>.npz.  Nothing from the reference is copied into the repo:
only its numeric outputs.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/2Haff/model")

from oracle import synth  # noqa: E402

# token positions whose residual-stream rows are saved after every block:
# window interiors, window edges (13|14), the padded last window (56..63)
TAP_POS = (0, 13, 14, 37, 56, 63)

CASES = {
    #  name      arch      n_frames n_prompts resize        original
    "tiny64": ("tiny64", 1, 2, (576, 1024), (270, 480)),
    "tiny80": ("tiny80", 2, 3, (1024, 683), (300, 200)),
    "vit_b_cfg1": ("vit_b", 1, 2, (576, 1024), (1080, 1920)),   # BASELINE.json configs[0]
    "vit_h_cfg2": ("vit_h", 2, 3, (576, 1024), (1080, 1920)),   # 2 frames of configs[1]
}


# "lite" cases: frames [first, first + n) of a config, outputs only (low-res logits as fp16, packed full-res mask
# bits, both heads, strided full-res samples) - no per-block taps, so eight ViT-H frames stay a few MB each
LITE_CASES = {
    #  name             arch    first n_frames n_prompts resize        original
    "vit_h_cfg2_rest": ("vit_h", 2, 6, 3, (576, 1024), (1080, 1920)),   # frames 2..7 of configs[1]
    "vit_l_lite": ("vit_l", 0, 1, 3, (576, 1024), (1080, 1920)),        # build_sam_vit_l, never a headline config
}


def build_reference(arch: synth.Arch):
    import segment_anything as sa
    from segment_anything.build_sam import _build_sam
    sam = _build_sam(encoder_embed_dim=arch.dim, encoder_depth=arch.depth,
                     encoder_num_heads=arch.heads,
                     encoder_global_attn_indexes=list(arch.global_idx),
                     is_sam_pretrained=None, checkpoint=None)
    sd = synth.make_state_dict(arch, seed=0)
    sam.load_state_dict(sd, strict=True)   # proves names + shapes match the reference
    return sam.eval()


def run_case(name: str):
    lite = name in LITE_CASES
    if lite:
        arch_name, first, n_frames, n_prompts, resize, original = LITE_CASES[name]
    else:
        arch_name, n_frames, n_prompts, resize, original = CASES[name]
        first = 0
    arch = synth.ARCHS[arch_name]
    torch.set_num_threads(os.cpu_count())
    sam = build_reference(arch)
    frames = synth.make_frames(first, n_frames, valid_rows=resize[0])
    out = {}
    taps = []
    hooks = [] if lite else [blk.register_forward_hook(lambda m, i, o: taps.append(o.detach().clone()))
                             for blk in sam.image_encoder.blocks]
    t0 = time.time()
    with torch.no_grad():
        # LISA.get_visual_embs runs B=1 per call (LISA.py:157-168)
        emb = torch.cat([sam.image_encoder(frames[i:i + 1]) for i in range(n_frames)], 0)
        for h in hooks:
            h.remove()
        pos = torch.tensor(TAP_POS)
        for bi in range(0 if lite else arch.depth):
            rows = torch.cat([taps[f * arch.depth + bi] for f in range(n_frames)], 0)
            out[f"tap_block{bi}"] = rows[:, pos][:, :, pos].numpy()          # (B,6,6,D)
        out["emb_slice"] = emb[:, ::16].numpy()                              # 16 of 256 channels
        out["emb_absmax"] = emb.abs().amax(dim=(1, 2, 3)).numpy()
        out["emb_mean"] = emb.mean(dim=(1, 2, 3)).double().numpy()
        out["dense_pe_slice"] = sam.prompt_encoder.get_dense_pe()[0, ::32].numpy()
        for i in range(n_frames):
            seg = synth.make_seg_embeds(first + i, n_prompts)
            sparse, dense = sam.prompt_encoder(points=None, boxes=None, masks=None,
                                               text_embeds=seg.unsqueeze(1))
            ml, il, tax = sam.mask_decoder_left(
                image_embeddings=emb[i:i + 1], image_pe=sam.prompt_encoder.get_dense_pe(),
                sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense, multimask_output=False)
            mr, ir = sam.mask_decoder_right(
                image_embeddings=emb[i:i + 1], image_pe=sam.prompt_encoder.get_dense_pe(),
                sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense, multimask_output=False)
            # full 4-vector iou head (the caller-visible slice is [:, 0:1])
            full = {}
            for side, dec in (("left", sam.mask_decoder_left), ("right", sam.mask_decoder_right)):
                res = dec.predict_masks(image_embeddings=emb[i:i + 1],
                                        image_pe=sam.prompt_encoder.get_dense_pe(),
                                        sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense)
                full[side] = res[1]
            pl = sam.postprocess_masks(ml, input_size=resize, original_size=original)[:, 0]
            pr = sam.postprocess_masks(mr, input_size=resize, original_size=original)[:, 0]
            out[f"f{i}_lowres_left"] = ml[:, 0].numpy().astype(np.float16 if arch_name in ("vit_h", "vit_l") else np.float32)
            out[f"f{i}_lowres_right"] = mr[:, 0].numpy().astype(np.float16 if arch_name in ("vit_h", "vit_l") else np.float32)
            out[f"f{i}_lowres_left_absmax"] = ml[:, 0].abs().amax(dim=(1, 2)).numpy()
            out[f"f{i}_lowres_right_absmax"] = mr[:, 0].abs().amax(dim=(1, 2)).numpy()
            out[f"f{i}_iou_left"] = full["left"].numpy()
            out[f"f{i}_iou_right"] = full["right"].numpy()
            out[f"f{i}_taxonomy"] = tax.numpy()
            out[f"f{i}_bits_left"] = np.packbits((pl > 0).numpy(), axis=-1)
            out[f"f{i}_bits_right"] = np.packbits((pr > 0).numpy(), axis=-1)
            # strided samples of the full-res fp32 logits (postprocess parity)
            out[f"f{i}_full_left_samples"] = pl[:, ::37, ::41].numpy()
            out[f"f{i}_full_right_samples"] = pr[:, ::37, ::41].numpy()
            out[f"f{i}_full_left_absmax"] = pl.abs().amax(dim=(1, 2)).numpy()
            out[f"f{i}_full_right_absmax"] = pr.abs().amax(dim=(1, 2)).numpy()
    out["meta"] = np.array([n_frames, n_prompts, resize[0], resize[1], original[0], original[1], first])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {time.time() - t0:.1f}s -> {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    for c in (sys.argv[1:] or list(CASES) + list(LITE_CASES)):
        run_case(c)
