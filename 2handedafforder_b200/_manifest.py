"""Names and shapes of `visual_model.state_dict()` (reference: the modules built by
2Haff/model/segment_anything/build_sam.py:59-121; listing in SURVEY.md Appendix C).

The facade keeps one real nn.Parameter / buffer per entry so HF `from_pretrained`,
`load_state_dict`, `.bfloat16()`, `.cuda()` and `named_parameters()` filters in
2Haff/train_ds.py:205,236-244 behave as with the reference modules.
"""
from __future__ import annotations

from typing import List, Tuple

GRID = 64
WINDOW = 14
OUT_CH = 256

Entry = Tuple[str, Tuple[int, ...], bool]  # (name, shape, is_buffer)


def _lin(out: List[Entry], p: str, o: int, i: int) -> None:
    out.append((p + ".weight", (o, i), False))
    out.append((p + ".bias", (o,), False))


def _ln(out: List[Entry], p: str, n: int) -> None:
    out.append((p + ".weight", (n,), False))
    out.append((p + ".bias", (n,), False))


def _attn(out: List[Entry], p: str, internal: int) -> None:
    for nm in ("q_proj", "k_proj", "v_proj"):
        _lin(out, f"{p}.{nm}", internal, OUT_CH)
    _lin(out, p + ".out_proj", OUT_CH, internal)


def manifest(dim: int, depth: int, heads: int, global_idx) -> List[Entry]:
    hd = dim // heads
    m: List[Entry] = []
    e = "image_encoder."
    m.append((e + "pos_embed", (1, GRID, GRID, dim), False))
    m.append((e + "patch_embed.proj.weight", (dim, 3, 16, 16), False))
    m.append((e + "patch_embed.proj.bias", (dim,), False))
    for i in range(depth):
        b = f"{e}blocks.{i}."
        S = GRID if i in global_idx else WINDOW
        _ln(m, b + "norm1", dim)
        m.append((b + "attn.rel_pos_h", (2 * S - 1, hd), False))
        m.append((b + "attn.rel_pos_w", (2 * S - 1, hd), False))
        _lin(m, b + "attn.qkv", 3 * dim, dim)
        _lin(m, b + "attn.proj", dim, dim)
        _ln(m, b + "norm2", dim)
        _lin(m, b + "mlp.lin1", 4 * dim, dim)
        _lin(m, b + "mlp.lin2", dim, 4 * dim)
    m.append((e + "neck.0.weight", (OUT_CH, dim, 1, 1), False))
    _ln(m, e + "neck.1", OUT_CH)
    m.append((e + "neck.2.weight", (OUT_CH, OUT_CH, 3, 3), False))
    _ln(m, e + "neck.3", OUT_CH)

    p = "prompt_encoder."
    m.append((p + "pe_layer.positional_encoding_gaussian_matrix", (2, OUT_CH // 2), True))
    for k in range(4):
        m.append((p + f"point_embeddings.{k}.weight", (1, OUT_CH), False))
    m.append((p + "not_a_point_embed.weight", (1, OUT_CH), False))
    m.append((p + "mask_downscaling.0.weight", (4, 1, 2, 2), False))
    m.append((p + "mask_downscaling.0.bias", (4,), False))
    _ln(m, p + "mask_downscaling.1", 4)
    m.append((p + "mask_downscaling.3.weight", (16, 4, 2, 2), False))
    m.append((p + "mask_downscaling.3.bias", (16,), False))
    _ln(m, p + "mask_downscaling.4", 16)
    m.append((p + "mask_downscaling.6.weight", (OUT_CH, 16, 1, 1), False))
    m.append((p + "mask_downscaling.6.bias", (OUT_CH,), False))
    m.append((p + "no_mask_embed.weight", (1, OUT_CH), False))

    for side in ("left", "right"):
        d = f"mask_decoder_{side}."
        m.append((d + "iou_token.weight", (1, OUT_CH), False))
        m.append((d + "mask_tokens.weight", (4, OUT_CH), False))
        for l in range(2):
            t = f"{d}transformer.layers.{l}."
            _attn(m, t + "self_attn", OUT_CH)
            _ln(m, t + "norm1", OUT_CH)
            _attn(m, t + "cross_attn_token_to_image", OUT_CH // 2)
            _ln(m, t + "norm2", OUT_CH)
            _lin(m, t + "mlp.lin1", 2048, OUT_CH)
            _lin(m, t + "mlp.lin2", OUT_CH, 2048)
            _ln(m, t + "norm3", OUT_CH)
            _ln(m, t + "norm4", OUT_CH)
            _attn(m, t + "cross_attn_image_to_token", OUT_CH // 2)
        _attn(m, d + "transformer.final_attn_token_to_image", OUT_CH // 2)
        _ln(m, d + "transformer.norm_final_attn", OUT_CH)
        m.append((d + "output_upscaling.0.weight", (256, 64, 2, 2), False))
        m.append((d + "output_upscaling.0.bias", (64,), False))
        _ln(m, d + "output_upscaling.1", 64)
        m.append((d + "output_upscaling.3.weight", (64, 32, 2, 2), False))
        m.append((d + "output_upscaling.3.bias", (32,), False))
        for k in range(4):
            h = f"{d}output_hypernetworks_mlps.{k}.layers."
            _lin(m, h + "0", 256, 256)
            _lin(m, h + "1", 256, 256)
            _lin(m, h + "2", 32, 256)
        h = d + "iou_prediction_head.layers."
        _lin(m, h + "0", 256, 256)
        _lin(m, h + "1", 256, 256)
        _lin(m, h + "2", 4, 256)
        if side == "left":
            h = d + "taxonomy_embed.layers."
            _lin(m, h + "0", 1024, 1024)
            _lin(m, h + "1", 1024, 1024)
            _lin(m, h + "2", 4, 1024)
    return m
