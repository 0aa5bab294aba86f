"""Seeded synthetic weights / frames / [SEG] embeddings for the SAM branch.

TEST INFRASTRUCTURE ONLY (see oracle/README.md): imported by tests/, bench.py's
cpu_baseline / --impl reference leg, __graft_entry__.smoke() and
tests/golden/make_golden.py.  The product package never imports this module.

The reference has no generator of its own (it loads sam_vit_h_4b8939.pth,
2Haff/README.md:72, which is not available offline), so this module draws a
state_dict with exactly the reference's key names and shapes (SURVEY.md
Appendix C; verified by `load_state_dict(strict=True)` into the reference's
`build_sam_vit_*` modules in tests/golden/make_golden.py).  Everything is
drawn from CPU `torch.Generator`s so the GPU box regenerates identical bytes.

Parity traps handled here: `pos_embed` and
`rel_pos_h/w` are zero-initialised in the reference
(2Haff/model/segment_anything/modeling/image_encoder.py:70-74,232-233); they
are drawn from N(0, 0.02) so those code paths are exercised.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch

IMG = 1024          # build_sam.py:68
PATCH = 16          # build_sam.py:69
GRID = IMG // PATCH  # 64
WINDOW = 14         # build_sam.py:84
OUT_CH = 256        # build_sam.py:67 prompt_embed_dim


@dataclass(frozen=True)
class Arch:
    name: str
    dim: int
    depth: int
    heads: int
    global_idx: Tuple[int, ...]
    arch_id: int  # C-ABI enum (include/b200sam.h)

    @property
    def head_dim(self) -> int:
        return self.dim // self.heads


# build_sam.py:15-48 hyper-parameter table, plus two tiny test-only shapes that
# exercise head_dim 64 and 80 in seconds on a CPU.
ARCHS: Dict[str, Arch] = {
    "vit_b": Arch("vit_b", 768, 12, 12, (2, 5, 8, 11), 0),
    "vit_l": Arch("vit_l", 1024, 24, 16, (5, 11, 17, 23), 1),
    "vit_h": Arch("vit_h", 1280, 32, 16, (7, 15, 23, 31), 2),
    "tiny64": Arch("tiny64", 256, 2, 4, (1,), 3),
    "tiny80": Arch("tiny80", 320, 3, 4, (2,), 4),
}


def _lin(g, out_f, in_f, prefix, sd, bias=True):
    bound = 1.0 / math.sqrt(in_f)
    sd[prefix + ".weight"] = (torch.rand(out_f, in_f, generator=g) * 2 - 1) * bound
    if bias:
        sd[prefix + ".bias"] = (torch.rand(out_f, generator=g) * 2 - 1) * bound


def _ln(g, n, prefix, sd):
    sd[prefix + ".weight"] = 1.0 + 0.1 * torch.randn(n, generator=g)
    sd[prefix + ".bias"] = 0.05 * torch.randn(n, generator=g)


def _attn(g, prefix, sd, dim, internal):
    for nm in ("q_proj", "k_proj", "v_proj"):
        _lin(g, internal, dim, f"{prefix}.{nm}", sd)
    _lin(g, dim, internal, f"{prefix}.out_proj", sd)


def make_state_dict(arch: Arch, seed: int = 0) -> Dict[str, torch.Tensor]:
    """fp32 state_dict with the reference's names/shapes."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    D, hd = arch.dim, arch.head_dim
    e = "image_encoder."
    sd[e + "pos_embed"] = 0.02 * torch.randn(1, GRID, GRID, D, generator=g)
    bound = 1.0 / math.sqrt(3 * PATCH * PATCH)
    sd[e + "patch_embed.proj.weight"] = (torch.rand(D, 3, PATCH, PATCH, generator=g) * 2 - 1) * bound
    sd[e + "patch_embed.proj.bias"] = (torch.rand(D, generator=g) * 2 - 1) * bound
    for i in range(arch.depth):
        b = f"{e}blocks.{i}."
        _ln(g, D, b + "norm1", sd)
        S = GRID if i in arch.global_idx else WINDOW
        sd[b + "attn.rel_pos_h"] = 0.02 * torch.randn(2 * S - 1, hd, generator=g)
        sd[b + "attn.rel_pos_w"] = 0.02 * torch.randn(2 * S - 1, hd, generator=g)
        _lin(g, 3 * D, D, b + "attn.qkv", sd)
        _lin(g, D, D, b + "attn.proj", sd)
        _ln(g, D, b + "norm2", sd)
        _lin(g, 4 * D, D, b + "mlp.lin1", sd)
        _lin(g, D, 4 * D, b + "mlp.lin2", sd)
    sd[e + "neck.0.weight"] = (torch.rand(OUT_CH, D, 1, 1, generator=g) * 2 - 1) / math.sqrt(D)
    _ln(g, OUT_CH, e + "neck.1", sd)
    sd[e + "neck.2.weight"] = (torch.rand(OUT_CH, OUT_CH, 3, 3, generator=g) * 2 - 1) / math.sqrt(OUT_CH * 9)
    _ln(g, OUT_CH, e + "neck.3", sd)

    p = "prompt_encoder."
    sd[p + "pe_layer.positional_encoding_gaussian_matrix"] = torch.randn(2, OUT_CH // 2, generator=g)
    for k in range(4):
        sd[p + f"point_embeddings.{k}.weight"] = torch.randn(1, OUT_CH, generator=g)
    sd[p + "not_a_point_embed.weight"] = torch.randn(1, OUT_CH, generator=g)
    # mask_downscaling is unused on the text-prompt path but must load (Appendix C)
    sd[p + "mask_downscaling.0.weight"] = 0.5 * torch.randn(4, 1, 2, 2, generator=g)
    sd[p + "mask_downscaling.0.bias"] = 0.1 * torch.randn(4, generator=g)
    _ln(g, 4, p + "mask_downscaling.1", sd)
    sd[p + "mask_downscaling.3.weight"] = 0.25 * torch.randn(16, 4, 2, 2, generator=g)
    sd[p + "mask_downscaling.3.bias"] = 0.1 * torch.randn(16, generator=g)
    _ln(g, 16, p + "mask_downscaling.4", sd)
    sd[p + "mask_downscaling.6.weight"] = 0.25 * torch.randn(OUT_CH, 16, 1, 1, generator=g)
    sd[p + "mask_downscaling.6.bias"] = 0.1 * torch.randn(OUT_CH, generator=g)
    sd[p + "no_mask_embed.weight"] = torch.randn(1, OUT_CH, generator=g)

    for side in ("left", "right"):
        d = f"mask_decoder_{side}."
        sd[d + "iou_token.weight"] = torch.randn(1, OUT_CH, generator=g)
        sd[d + "mask_tokens.weight"] = torch.randn(4, OUT_CH, generator=g)
        for l in range(2):
            t = f"{d}transformer.layers.{l}."
            _attn(g, t + "self_attn", sd, OUT_CH, OUT_CH)
            _ln(g, OUT_CH, t + "norm1", sd)
            _attn(g, t + "cross_attn_token_to_image", sd, OUT_CH, OUT_CH // 2)
            _ln(g, OUT_CH, t + "norm2", sd)
            _lin(g, 2048, OUT_CH, t + "mlp.lin1", sd)
            _lin(g, OUT_CH, 2048, t + "mlp.lin2", sd)
            _ln(g, OUT_CH, t + "norm3", sd)
            _ln(g, OUT_CH, t + "norm4", sd)
            _attn(g, t + "cross_attn_image_to_token", sd, OUT_CH, OUT_CH // 2)
        _attn(g, d + "transformer.final_attn_token_to_image", sd, OUT_CH, OUT_CH // 2)
        _ln(g, OUT_CH, d + "transformer.norm_final_attn", sd)
        # ConvTranspose2d weights are (in, out, kH, kW)
        sd[d + "output_upscaling.0.weight"] = (torch.rand(256, 64, 2, 2, generator=g) * 2 - 1) / math.sqrt(64 * 4)
        sd[d + "output_upscaling.0.bias"] = (torch.rand(64, generator=g) * 2 - 1) / math.sqrt(64 * 4)
        _ln(g, 64, d + "output_upscaling.1", sd)
        sd[d + "output_upscaling.3.weight"] = (torch.rand(64, 32, 2, 2, generator=g) * 2 - 1) / math.sqrt(32 * 4)
        sd[d + "output_upscaling.3.bias"] = (torch.rand(32, generator=g) * 2 - 1) / math.sqrt(32 * 4)
        for m in range(4):
            h = f"{d}output_hypernetworks_mlps.{m}.layers."
            _lin(g, 256, 256, h + "0", sd)
            _lin(g, 256, 256, h + "1", sd)
            _lin(g, 32, 256, h + "2", sd)
        h = d + "iou_prediction_head.layers."
        _lin(g, 256, 256, h + "0", sd)
        _lin(g, 256, 256, h + "1", sd)
        _lin(g, 4, 256, h + "2", sd)
        if side == "left":  # build_sam.py:103,116 taxonomy_on only on the left decoder
            h = d + "taxonomy_embed.layers."
            _lin(g, 1024, 1024, h + "0", sd)
            _lin(g, 1024, 1024, h + "1", sd)
            _lin(g, 4, 1024, h + "2", sd)
    return sd


def make_frame(index: int, valid_rows: int = 576) -> torch.Tensor:
    """One normalised, zero-padded 1024^2 frame, seed 10+index.

    N(0,1) stands in for mean/std-normalised pixels; rows >= valid_rows are
    zero, mimicking the bottom padding a 16:9 frame gets in
    2Haff/inference.py:91-105.
    """
    g = torch.Generator().manual_seed(10 + index)
    x = torch.randn(3, IMG, IMG, generator=g)
    x[:, valid_rows:, :] = 0.0
    return x


def make_frames(start: int, count: int, valid_rows: int = 576) -> torch.Tensor:
    return torch.stack([make_frame(start + i, valid_rows) for i in range(count)], 0)


def make_seg_embeds(index: int, n_prompts: int) -> torch.Tensor:
    """(n_prompts, 256) projected [SEG] hidden states for frame `index`."""
    g = torch.Generator().manual_seed(100003 + index)
    return torch.randn(n_prompts, OUT_CH, generator=g)


def flops_per_frame(arch: Arch, n_prompts: int = 3) -> float:
    """Algorithmic FLOPs per frame (multiply-add = 2, unpadded tokens)."""
    D, T = arch.dim, GRID * GRID
    hd = arch.head_dim
    f = 2.0 * T * (3 * PATCH * PATCH) * D                       # patch embed
    per_block_gemm = 2.0 * T * D * (3 * D + D + 8 * D)
    f += arch.depth * per_block_gemm
    n_glob = len(arch.global_idx)
    n_win = arch.depth - n_glob
    f += n_win * (2.0 * 2 * T * (WINDOW * WINDOW) * D)           # QK^T + PV, 196 keys/query
    f += n_glob * (2.0 * 2 * T * T * D)
    f += n_win * (2.0 * 2 * T * WINDOW * D) + n_glob * (2.0 * 2 * T * GRID * D)  # rel-pos
    f += 2.0 * T * D * OUT_CH + 2.0 * T * OUT_CH * OUT_CH * 9    # neck
    f += n_prompts * 2 * 3.606e9 + n_prompts * 0.004e9          # decoders
    return f


import numpy as np  # noqa: E402


def synth_image(H, W, seed):
    """(H,W,3) uint8 RGB test frame for the preprocessing row.  Smooth structure + noise so both the antialias window and the rounding paths are exercised."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W]
    base = 127 + 100 * np.sin(xx / 17.0 + seed) * np.cos(yy / 11.0)
    img = base[..., None] + rng.normal(0, 40, (H, W, 3))
    return np.clip(img, 0, 255).astype(np.uint8)


# --------------------------------------------------------------------------- training case
TRAIN_CASE = {
    # two frames with 2 and 1 [SEG] prompts, different aspect ratios, small label resolutions
    "prompts": (2, 1),
    "resize": ((576, 1024), (1024, 683)),
    "original": ((135, 240), (150, 100)),
    "taxonomies": ((0.0, 0.0, 1.0, 0.0), (1.0, 0.0, 0.0, 0.0)),   # frame 0: bimanual, frame 1: left hand only
}


def make_train_inputs(seed: int = 7):
    """Image embeddings (F,256,64,64) (stand-ins for the frozen encoder's output), ground-truth masks per frame and
    side, one-hot taxonomies.  The [SEG] embeddings are make_seg_embeds(i, n)."""
    g = torch.Generator().manual_seed(seed)
    n_frames = len(TRAIN_CASE["prompts"])
    emb = 0.5 * torch.randn(n_frames, OUT_CH, GRID, GRID, generator=g)
    gt_left, gt_right = [], []
    for i, (n, (h, w)) in enumerate(zip(TRAIN_CASE["prompts"], TRAIN_CASE["original"])):
        yy = torch.arange(h, dtype=torch.float32)[:, None] / h
        xx = torch.arange(w, dtype=torch.float32)[None, :] / w
        c = torch.rand(n, 2, 2, generator=g)
        gt_left.append((((yy - c[:, 0, 0, None, None]) ** 2 + (xx - c[:, 0, 1, None, None]) ** 2) < 0.06).float())
        gt_right.append((((yy - c[:, 1, 0, None, None]) ** 2 + (xx - c[:, 1, 1, None, None]) ** 2) < 0.04).float())
    tax = torch.tensor(TRAIN_CASE["taxonomies"], dtype=torch.float32)
    segs = [make_seg_embeds(i, n) for i, n in enumerate(TRAIN_CASE["prompts"])]
    return emb, segs, gt_left, gt_right, tax
