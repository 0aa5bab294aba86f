"""CPU oracle for the SAM segmentation branch of 2HandedAfforder.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / `--impl reference` leg may import this module; the product
package (2handedafforder_b200/) never does and raises if its CUDA library is
missing.

This is a from-the-math restatement (plain fp32 tensor algebra: matmul, exp,
erf, explicit gathers) of the reference's algorithm, NOT a call-through to its
nn.Modules.  Every function cites the reference file:line it follows
(paths relative to 2Haff/model/segment_anything/).  Convolutions are written
as the GEMMs the CUDA kernels use (im2col order, pixel-shuffle order), so the
oracle also pins those index maps.

Parity status: PINNED against outputs of the reference itself, run in the
build container by tests/golden/make_golden.py (the reference ships no tests
or golden vectors: SURVEY.md section 4).  tests/test_oracle_golden.py checks
this module against those committed fixtures on every CPU run.

`Precision` emulates the GPU numerics recipe on the CPU (operand rounding of
the encoder GEMMs to bf16, of the decoder GEMMs to tf32) so the accuracy
budget of DESIGN.md can be checked without a GPU.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .synth import GRID, IMG, OUT_CH, PATCH, WINDOW, Arch

W = Dict[str, torch.Tensor]


@dataclass(frozen=True)
class Precision:
    enc: str = "fp32"   # "fp32" | "bf16"  (GEMM/attention operand rounding, fp32 accumulate)
    dec: str = "fp32"   # "fp32" | "tf32" | "bf16"


FP32 = Precision()
GPU_RECIPE = Precision(enc="bf16", dec="tf32")


def _rnd(x: torch.Tensor, mode: str) -> torch.Tensor:
    if mode == "fp32":
        return x
    if mode == "bf16":
        return x.to(torch.bfloat16).to(torch.float32)
    if mode == "tf32":  # tensor cores read the top 19 bits of an fp32 operand
        return (x.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)
    raise ValueError(mode)


def _mm(a: torch.Tensor, w: torch.Tensor, mode: str) -> torch.Tensor:
    """a (.., K) times w (N, K) transposed, operands rounded per `mode`."""
    return _rnd(a, mode) @ _rnd(w, mode).t()


def layer_norm(x, g, b, eps):
    """nn.LayerNorm over the last dim (biased variance)."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * g + b


def gelu_erf(x):
    """nn.GELU() default = exact erf form (common.py:18, image_encoder.py:30)."""
    return 0.5 * x * (1.0 + torch.erf(x * (1.0 / math.sqrt(2.0))))


# --------------------------------------------------------------------------- encoder

def patch_embed(w: W, frames: torch.Tensor, pm: str) -> torch.Tensor:
    """PatchEmbed Conv2d(3->D, k16, s16) + bias as a GEMM (image_encoder.py:395-426).

    im2col column order is (c, ky, kx), matching `weight.reshape(D, 768)`.
    Returns tokens (B, 64, 64, D) = the reference's permute(0, 2, 3, 1).
    """
    B = frames.shape[0]
    a = frames.reshape(B, 3, GRID, PATCH, GRID, PATCH).permute(0, 2, 4, 1, 3, 5)
    a = a.reshape(B, GRID, GRID, 3 * PATCH * PATCH)
    wt = w["image_encoder.patch_embed.proj.weight"].reshape(-1, 3 * PATCH * PATCH)
    return _mm(a, wt, pm) + w["image_encoder.patch_embed.proj.bias"]


def window_partition(x: torch.Tensor) -> torch.Tensor:
    """image_encoder.py:263-288: zero-pad 64->70 bottom/right, cut 25 windows of 14x14."""
    B, H, Wd, C = x.shape
    pad = (WINDOW - H % WINDOW) % WINDOW
    xp = torch.zeros(B, H + pad, Wd + pad, C, dtype=x.dtype)
    xp[:, :H, :Wd] = x
    n = (H + pad) // WINDOW
    xp = xp.reshape(B, n, WINDOW, n, WINDOW, C).permute(0, 1, 3, 2, 4, 5)
    return xp.reshape(B * n * n, WINDOW, WINDOW, C)


def window_unpartition(xw: torch.Tensor, B: int) -> torch.Tensor:
    """image_encoder.py:291-318: inverse of window_partition, dropping the pad."""
    n = (GRID + WINDOW - 1) // WINDOW
    C = xw.shape[-1]
    x = xw.reshape(B, n, n, WINDOW, WINDOW, C).permute(0, 1, 3, 2, 4, 5)
    return x.reshape(B, n * WINDOW, n * WINDOW, C)[:, :GRID, :GRID]


def vit_attention(w: W, pre: str, x: torch.Tensor, heads: int, pm: str) -> torch.Tensor:
    """Attention.forward with decomposed rel-pos (image_encoder.py:235-260, 321-392).

    x: (B', S, S, D).  No interpolation of the rel-pos tables is needed at
    1024^2 (table length is exactly 2S-1, image_encoder.py:333-335).
    The score uses q*scale, the bias uses the UNSCALED q (:244-249).
    """
    Bp, S, _, D = x.shape
    hd = D // heads
    qkv = _mm(x.reshape(Bp, S * S, D), w[pre + "qkv.weight"], pm) + w[pre + "qkv.bias"]
    qkv = _rnd(qkv, pm)  # the GPU path stores q/k/v as bf16
    qkv = qkv.reshape(Bp, S * S, 3, heads, hd).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]                      # (B', heads, T, hd)
    attn = (q * (hd ** -0.5)) @ k.transpose(-2, -1)
    idx = torch.arange(S)[:, None] - torch.arange(S)[None, :] + (S - 1)   # (q, k)
    Rh = _rnd(w[pre + "rel_pos_h"], pm)[idx]               # (S, S, hd)
    Rw = _rnd(w[pre + "rel_pos_w"], pm)[idx]
    rq = q.reshape(Bp, heads, S, S, hd)
    rel_h = torch.einsum("bnhwc,hkc->bnhwk", rq, Rh)
    rel_w = torch.einsum("bnhwc,wkc->bnhwk", rq, Rw)
    attn = attn.reshape(Bp, heads, S, S, S, S) + rel_h[..., :, None] + rel_w[..., None, :]
    attn = torch.softmax(attn.reshape(Bp, heads, S * S, S * S), dim=-1)
    o = _rnd(attn, pm) @ v                                 # P is fed to the PV MMA as bf16
    o = o.permute(0, 2, 1, 3).reshape(Bp, S, S, D)
    return _mm(o, w[pre + "proj.weight"], pm) + w[pre + "proj.bias"]


def encoder_block(w: W, arch: Arch, i: int, x: torch.Tensor, pm: str) -> torch.Tensor:
    """Block.forward (image_encoder.py:177-193); LN eps 1e-6 (build_sam.py:77)."""
    pre = f"image_encoder.blocks.{i}."
    B = x.shape[0]
    y = layer_norm(x, w[pre + "norm1.weight"], w[pre + "norm1.bias"], 1e-6)
    if i in arch.global_idx:
        y = vit_attention(w, pre + "attn.", y, arch.heads, pm)
    else:
        # pad AFTER the norm: pad tokens are zeros -> qkv = bias, and they are
        # real keys in the softmax (Appendix A1).
        y = window_unpartition(vit_attention(w, pre + "attn.", window_partition(y), arch.heads, pm), B)
    x = x + y
    y = layer_norm(x, w[pre + "norm2.weight"], w[pre + "norm2.bias"], 1e-6)
    h = gelu_erf(_mm(y, w[pre + "mlp.lin1.weight"], pm) + w[pre + "mlp.lin1.bias"])
    return x + _mm(h, w[pre + "mlp.lin2.weight"], pm) + w[pre + "mlp.lin2.bias"]


def neck(w: W, x: torch.Tensor, pm: str) -> torch.Tensor:
    """neck (image_encoder.py:92-108): conv1x1 -> LayerNorm2d -> conv3x3 pad1 -> LayerNorm2d.

    Token-major in, token-major (B, 64, 64, 256) out.  LayerNorm2d over the
    channel dim of one pixel (common.py:38-43) == row LayerNorm here.  The 3x3
    conv is a 9-tap implicit GEMM with K order (tap = dy*3+dx, c).
    """
    e = "image_encoder.neck."
    B = x.shape[0]
    z = _mm(x, w[e + "0.weight"].reshape(OUT_CH, -1), pm)
    z = layer_norm(z, w[e + "1.weight"], w[e + "1.bias"], 1e-6)
    zp = torch.zeros(B, GRID + 2, GRID + 2, OUT_CH)
    zp[:, 1:-1, 1:-1] = z
    cols = [zp[:, dy:dy + GRID, dx:dx + GRID] for dy in range(3) for dx in range(3)]
    a = torch.cat(cols, dim=-1)                                    # (B,64,64,9*256)
    w2 = w[e + "2.weight"].permute(0, 2, 3, 1).reshape(OUT_CH, 9 * OUT_CH)  # (out, dy, dx, c)
    z = _mm(a, w2, pm)
    return layer_norm(z, w[e + "3.weight"], w[e + "3.bias"], 1e-6)


def encode(w: W, arch: Arch, frames: torch.Tensor, prec: Precision = FP32,
           taps: Optional[dict] = None) -> torch.Tensor:
    """ImageEncoderViT.forward (image_encoder.py:110-125).  Returns NCHW (B,256,64,64)."""
    pm = prec.enc
    frames = _rnd(frames.float(), pm)
    x = patch_embed(w, frames, pm) + w["image_encoder.pos_embed"]
    if taps is not None:
        taps["patch"] = x.clone()
    for i in range(arch.depth):
        x = encoder_block(w, arch, i, x, pm)
        if taps is not None:
            taps[f"block{i}"] = x.clone()
    return neck(w, x, pm).permute(0, 3, 1, 2).contiguous()


# --------------------------------------------------------------------------- prompt encoder

def dense_pe(w: W) -> torch.Tensor:
    """PromptEncoder.get_dense_pe (prompt_encoder.py:67-76, 203-229) -> (1,256,64,64)."""
    G = w["prompt_encoder.pe_layer.positional_encoding_gaussian_matrix"]
    c = (torch.arange(GRID, dtype=torch.float32, device=G.device) + 0.5) / GRID
    xy = torch.stack([c[None, :].expand(GRID, GRID), c[:, None].expand(GRID, GRID)], dim=-1)
    ang = (2.0 * math.pi) * ((2.0 * xy - 1.0) @ G)
    return torch.cat([torch.sin(ang), torch.cos(ang)], dim=-1).permute(2, 0, 1)[None]


def prompt_encode(w: W, seg_embeds: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """PromptEncoder.forward, text path (prompt_encoder.py:140-186).

    seg_embeds (N, 256) -> sparse (N,1,256), dense (N,256,64,64) (= no_mask_embed broadcast).
    """
    N = seg_embeds.shape[0]
    dense = w["prompt_encoder.no_mask_embed.weight"].reshape(1, OUT_CH, 1, 1).expand(N, OUT_CH, GRID, GRID)
    return seg_embeds.float()[:, None, :], dense


# --------------------------------------------------------------------------- decoder

def _attn(w: W, pre: str, q, k, v, dm: str, heads: int = 8):
    """transformer.Attention.forward (transformer.py:220-242)."""
    q = _mm(q, w[pre + "q_proj.weight"], dm) + w[pre + "q_proj.bias"]
    k = _mm(k, w[pre + "k_proj.weight"], dm) + w[pre + "k_proj.bias"]
    v = _mm(v, w[pre + "v_proj.weight"], dm) + w[pre + "v_proj.bias"]
    N, nq, C = q.shape
    c = C // heads
    qh = q.reshape(N, nq, heads, c).transpose(1, 2)
    kh = k.reshape(N, k.shape[1], heads, c).transpose(1, 2)
    vh = v.reshape(N, v.shape[1], heads, c).transpose(1, 2)
    a = torch.softmax((qh @ kh.transpose(-2, -1)) / math.sqrt(c), dim=-1)
    o = (a @ vh).transpose(1, 2).reshape(N, nq, C)
    return _mm(o, w[pre + "out_proj.weight"], dm) + w[pre + "out_proj.bias"]


def _mlp3(w: W, pre: str, x, dm: str):
    """mask_decoder.MLP with 3 layers, ReLU between (mask_decoder.py:183-205)."""
    for i in range(3):
        x = _mm(x, w[f"{pre}layers.{i}.weight"], dm) + w[f"{pre}layers.{i}.bias"]
        if i < 2:
            x = torch.relu(x)
    return x


def two_way_transformer(w: W, pre: str, src_tok, pe_tok, tokens, dm: str):
    """TwoWayTransformer.forward (transformer.py:62-106, 151-182).  LN eps 1e-5."""
    def ln(x, name):
        return layer_norm(x, w[pre + name + ".weight"], w[pre + name + ".bias"], 1e-5)

    queries, keys = tokens, src_tok
    for l in range(2):
        p = f"{pre}layers.{l}."
        if l == 0:   # skip_first_layer_pe: REPLACES queries (transformer.py:155-156)
            queries = _attn(w, p + "self_attn.", queries, queries, queries, dm)
        else:
            q = queries + tokens
            queries = queries + _attn(w, p + "self_attn.", q, q, queries, dm)
        queries = ln(queries, f"layers.{l}.norm1")
        q = queries + tokens
        k = keys + pe_tok
        queries = ln(queries + _attn(w, p + "cross_attn_token_to_image.", q, k, keys, dm), f"layers.{l}.norm2")
        h = torch.relu(_mm(queries, w[p + "mlp.lin1.weight"], dm) + w[p + "mlp.lin1.bias"])
        queries = ln(queries + _mm(h, w[p + "mlp.lin2.weight"], dm) + w[p + "mlp.lin2.bias"], f"layers.{l}.norm3")
        q = queries + tokens
        k = keys + pe_tok
        keys = ln(keys + _attn(w, p + "cross_attn_image_to_token.", k, q, queries, dm), f"layers.{l}.norm4")
    q = queries + tokens
    k = keys + pe_tok
    queries = ln(queries + _attn(w, pre + "final_attn_token_to_image.", q, k, keys, dm), "norm_final_attn")
    return queries, keys


def decode(w: W, side: str, emb: torch.Tensor, seg_embeds: torch.Tensor,
           prec: Precision = FP32) -> Dict[str, torch.Tensor]:
    """MaskDecoder.forward + predict_masks for ONE frame (mask_decoder.py:79-178).

    emb (1,256,64,64) or (256,64,64); seg_embeds (N,256).
    Returns masks_all (N,4,256,256), iou_all (N,4), and the caller-visible
    slices masks (N,1,256,256), iou (N,1) for multimask_output=False
    (mask_decoder.py:110-115), plus taxonomy (N,4) for side == "left".
    """
    dm = prec.dec
    d = f"mask_decoder_{side}."
    N = seg_embeds.shape[0]
    sparse, dense = prompt_encode(w, seg_embeds)
    out_tok = torch.cat([w[d + "iou_token.weight"], w[d + "mask_tokens.weight"]], 0)
    tokens = torch.cat([out_tok[None].expand(N, -1, -1), sparse], dim=1)          # (N,6,256)
    src = emb.reshape(1, OUT_CH, GRID, GRID).float() + dense                       # (N,256,64,64)
    src_tok = src.flatten(2).permute(0, 2, 1)                                      # (N,4096,256)
    pe_tok = dense_pe(w).flatten(2).permute(0, 2, 1).expand(N, -1, -1)
    hs, keys = two_way_transformer(w, d + "transformer.", src_tok, pe_tok, tokens, dm)
    iou_tok, mask_toks = hs[:, 0], hs[:, 1:5]

    # output_upscaling (mask_decoder.py:54-64): ConvT(k2,s2) == GEMM + pixel shuffle.
    w0 = w[d + "output_upscaling.0.weight"]                                        # (256, 64, 2, 2) = (in,out,ky,kx)
    u = _mm(keys, w0.permute(2, 3, 1, 0).reshape(4 * 64, 256), dm)                 # cols (ky,kx,out)
    u = u.reshape(N, GRID, GRID, 2, 2, 64) + w[d + "output_upscaling.0.bias"]
    u = u.permute(0, 1, 3, 2, 4, 5).reshape(N, 128, 128, 64)
    u = gelu_erf(layer_norm(u, w[d + "output_upscaling.1.weight"], w[d + "output_upscaling.1.bias"], 1e-6))
    w3 = w[d + "output_upscaling.3.weight"]                                        # (64, 32, 2, 2)
    u2 = _mm(u, w3.permute(2, 3, 1, 0).reshape(4 * 32, 64), dm)
    u2 = u2.reshape(N, 128, 128, 2, 2, 32) + w[d + "output_upscaling.3.bias"]
    u2 = gelu_erf(u2.permute(0, 1, 3, 2, 4, 5).reshape(N, 256, 256, 32))

    hyper = torch.stack([_mlp3(w, f"{d}output_hypernetworks_mlps.{m}.", mask_toks[:, m], dm)
                         for m in range(4)], dim=1)                                # (N,4,32)
    masks = torch.einsum("nmc,nyxc->nmyx", hyper, u2)                              # (N,4,256,256)
    iou = _mlp3(w, d + "iou_prediction_head.", iou_tok, dm)
    out = {"masks_all": masks, "iou_all": iou, "masks": masks[:, 0:1], "iou": iou[:, 0:1]}
    if side == "left":
        out["taxonomy"] = torch.softmax(_mlp3(w, d + "taxonomy_embed.", mask_toks.flatten(1), dm), dim=-1)
    return out


# --------------------------------------------------------------------------- postprocess

def _bilinear_axis(n_in: int, n_out: int, device=None):
    """ATen upsample_bilinear2d source index, align_corners=False (sam.py:177-189).

    src = max(0, (dst + 0.5) * in/out - 0.5); i0 = floor; i1 = min(i0+1, in-1); lam = src - i0.
    """
    scale = torch.tensor(n_in / n_out, dtype=torch.float32, device=device)
    dst = torch.arange(n_out, dtype=torch.float32, device=device)
    src = torch.clamp(scale * (dst + 0.5) - 0.5, min=0.0)
    i0 = src.floor().long().clamp(max=n_in - 1)
    i1 = torch.clamp(i0 + 1, max=n_in - 1)
    lam = src - i0.float()
    return i0, i1, lam


def bilinear_resize(x: torch.Tensor, out_hw: Tuple[int, int]) -> torch.Tensor:
    """F.interpolate(x, out_hw, mode='bilinear', align_corners=False) on (..., H, W)."""
    H, Wd = x.shape[-2:]
    y0, y1, ly = _bilinear_axis(H, int(out_hw[0]), x.device)
    x0, x1, lx = _bilinear_axis(Wd, int(out_hw[1]), x.device)
    top, bot = x[..., y0, :], x[..., y1, :]
    def lerp_x(r):
        return (1.0 - lx) * r[..., x0] + lx * r[..., x1]
    return (1.0 - ly)[:, None] * lerp_x(top) + ly[:, None] * lerp_x(bot)


def postprocess_masks(lowres: torch.Tensor, input_size, original_size) -> torch.Tensor:
    """Sam.postprocess_masks (sam.py:155-189): two SEQUENTIAL bilinear resamplings
    with a crop in between (Appendix A10).  Returns fp32 (N,C,H,W)."""
    m = bilinear_resize(lowres.float(), (IMG, IMG))
    m = m[..., : int(input_size[0]), : int(input_size[1])]
    return bilinear_resize(m, (int(original_size[0]), int(original_size[1])))


# --------------------------------------------------------------------------- whole path

def segment(w: W, arch: Arch, frames: torch.Tensor, seg_embeds: Sequence[torch.Tensor],
            resize_list, original_size_list, prec: Precision = FP32,
            emb: Optional[torch.Tensor] = None) -> List[Dict[str, torch.Tensor]]:
    """The path LISA.evaluate drives per image (2Haff/model/LISA.py:487-532)."""
    if emb is None:
        emb = encode(w, arch, frames, prec)
    outs = []
    for i in range(emb.shape[0]):
        L = decode(w, "left", emb[i], seg_embeds[i], prec)
        R = decode(w, "right", emb[i], seg_embeds[i], prec)
        outs.append({
            "lowres_left": L["masks"][:, 0], "lowres_right": R["masks"][:, 0],
            "iou_left": L["iou_all"], "iou_right": R["iou_all"], "taxonomy": L["taxonomy"],
            "masks_left": postprocess_masks(L["masks"], resize_list[i], original_size_list[i])[:, 0],
            "masks_right": postprocess_masks(R["masks"], resize_list[i], original_size_list[i])[:, 0],
        })
    return outs


# --------------------------------------------------------------------------- metrics (SURVEY.md 8d pass criteria)

def logit_rel_err(a: torch.Tensor, ref: torch.Tensor) -> float:
    """max|a-ref| / max|ref| per mask (pointwise relative error is meaningless at zero crossings)."""
    return float((a.float() - ref.float()).abs().max() / ref.float().abs().max().clamp_min(1e-30))


def mask_iou(a: torch.Tensor, ref: torch.Tensor, thr: float = 0.0) -> float:
    A, R = a > thr, ref > thr
    union = (A | R).sum().item()
    return 1.0 if union == 0 else (A & R).sum().item() / union
