"""Constructors with the reference's names and hyper-parameter table
(2Haff/model/segment_anything/build_sam.py:15-139), returning the B200-native `visual_model`.

`checkpoint` + `is_sam_pretrained` follow the reference: a raw SAM checkpoint has a single `mask_decoder.*`,
which is duplicated into `mask_decoder_left.*` and `mask_decoder_right.*` (build_sam.py:125-136); a 2Haff
checkpoint is loaded as is, `strict=False` in both cases.
"""
from __future__ import annotations

import torch

from .sam import B200Sam


def _build_sam(encoder_embed_dim, encoder_depth, encoder_num_heads, encoder_global_attn_indexes,
               checkpoint=None, is_sam_pretrained=None, **kw) -> B200Sam:
    sam = B200Sam(encoder_embed_dim, encoder_depth, encoder_num_heads, encoder_global_attn_indexes, **kw)
    sam.eval()
    if checkpoint is not None:
        with open(checkpoint, "rb") as f:
            state_dict = torch.load(f, map_location="cpu")
        if is_sam_pretrained:
            remapped = {}
            for key, value in state_dict.items():
                if key.startswith("mask_decoder"):
                    remapped[key.replace("mask_decoder", "mask_decoder_left", 1)] = value
                    remapped[key.replace("mask_decoder", "mask_decoder_right", 1)] = value
                else:
                    remapped[key] = value
            state_dict = remapped
        sam.load_state_dict(state_dict, strict=False)
    return sam


def build_sam_vit_h(checkpoint=None, is_sam_pretrained=None, **kw):
    return _build_sam(1280, 32, 16, [7, 15, 23, 31], checkpoint, is_sam_pretrained, **kw)


build_sam = build_sam_vit_h


def build_sam_vit_l(checkpoint=None, is_sam_pretrained=None, **kw):
    return _build_sam(1024, 24, 16, [5, 11, 17, 23], checkpoint, is_sam_pretrained, **kw)


def build_sam_vit_b(checkpoint=None, is_sam_pretrained=None, **kw):
    return _build_sam(768, 12, 12, [2, 5, 8, 11], checkpoint, is_sam_pretrained, **kw)


sam_model_registry = {
    "default": build_sam_vit_h,
    "vit_h": build_sam_vit_h,
    "vit_l": build_sam_vit_l,
    "vit_b": build_sam_vit_b,
}
