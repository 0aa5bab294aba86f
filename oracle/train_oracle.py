"""CPU oracle for the training forward / backward of the SAM branch.

TEST INFRASTRUCTURE ONLY (same rule as oracle/sam_oracle.py: only tests/, smoke() and bench.py's CPU legs import it).

The forward is the fp32 restatement of oracle/sam_oracle.py (decode + postprocess_masks, plain tensor algebra, already
pinned against the reference's inference outputs) plus the two mask losses and their per-image composition restated
here from 2Haff/model/LISA.py; the backward is torch.autograd over that restatement on the CPU - a floating-point
reference, as the task allows for floating-point kernels.

Parity status: PINNED.  tests/golden/make_golden_train.py runs the UNMODIFIED reference modules (MaskDecoder,
Sam.postprocess_masks) and the reference's own dice_loss / sigmoid_ce_loss (their source is loaded from
/root/reference at generation time, never copied) under autograd and stores losses, outputs, the full gradient of the
[SEG] embeddings and digests of every parameter gradient in tests/golden/train_tiny.npz;
tests/test_oracle_golden.py checks this module against that file on every CPU run.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch

from . import sam_oracle as O

W = Dict[str, torch.Tensor]


def dice_loss(inputs: torch.Tensor, targets: torch.Tensor, num_masks: float, scale: float = 1000.0,
              eps: float = 1e-6) -> torch.Tensor:
    """LISA.py:16-39.  inputs / targets (N, H, W); sigmoid, per-mask sums scaled by 1/scale, 1 - (2ab + eps)/(a + b + eps),
    summed over the masks and divided by num_masks + 1e-8."""
    s = torch.sigmoid(inputs).flatten(1, 2)
    t = targets.flatten(1, 2)
    num = 2.0 * (s / scale * t).sum(-1)
    den = (s / scale).sum(-1) + (t / scale).sum(-1)
    return (1.0 - (num + eps) / (den + eps)).sum() / (num_masks + 1e-8)


def sigmoid_ce_loss(inputs: torch.Tensor, targets: torch.Tensor, num_masks: float) -> torch.Tensor:
    """LISA.py:42-59.  binary_cross_entropy_with_logits written out: max(x, 0) - x t + log(1 + exp(-|x|)); mean over
    the pixels of each mask, summed over the masks, divided by num_masks + 1e-8."""
    x, t = inputs, targets
    l = torch.clamp(x, min=0) - x * t + torch.log1p(torch.exp(-x.abs()))
    return l.flatten(1, 2).mean(1).sum() / (num_masks + 1e-8)


def sam_losses(pred_left: Sequence[torch.Tensor], pred_right: Sequence[torch.Tensor], gt_left: Sequence[torch.Tensor],
               gt_right: Sequence[torch.Tensor], gt_taxonomies: torch.Tensor, bce_w: float = 2.0,
               dice_w: float = 0.5) -> Dict[str, torch.Tensor]:
    """The mask part of LISA.model_forward's loss (LISA.py:346-420): the taxonomy gate multiplies the logits
    (:357-358), per image the two losses are multiplied back by the number of masks (:389-407), and the totals are
    divided by num_masks + 1e-8 with the bce / dice weights (:416-418).  gt_taxonomies: (F, 4) one-hot rows
    [left, right, both-a, both-b]."""
    wl = gt_taxonomies[:, 0] + gt_taxonomies[:, 2] + gt_taxonomies[:, 3]
    wr = gt_taxonomies[:, 1] + gt_taxonomies[:, 2] + gt_taxonomies[:, 3]
    bl = br = dl = dr = 0.0
    n = 0
    for i in range(len(pred_left)):
        pl, pr = wl[i] * pred_left[i], wr[i] * pred_right[i]
        k = gt_left[i].shape[0]
        bl = bl + sigmoid_ce_loss(pl, gt_left[i], k) * k
        dl = dl + dice_loss(pl, gt_left[i], k) * k
        br = br + sigmoid_ce_loss(pr, gt_right[i], k) * k
        dr = dr + dice_loss(pr, gt_right[i], k) * k
        n += k
    bce = bce_w * bl / (n + 1e-8) + bce_w * br / (n + 1e-8)
    dice = dice_w * dl / (n + 1e-8) + dice_w * dr / (n + 1e-8)
    return {"mask_bce_loss": bce, "mask_dice_loss": dice, "mask_loss": bce + dice}


def train_step(w: W, emb: torch.Tensor, seg_embeds: Sequence[torch.Tensor], resize_list, original_list,
               gt_left: Sequence[torch.Tensor], gt_right: Sequence[torch.Tensor], gt_taxonomies: torch.Tensor,
               tax_weight: float = 1.0) -> Dict[str, object]:
    """The SAM branch of one training step as LISA.model_forward runs it (LISA.py:284-420) in fp32 on the CPU:
    per image prompt encoder -> both decoders -> postprocess_masks (:284-325), the mask losses (:346-418) and a
    cross-entropy of the taxonomy head against the one-hot ground truth (:411-415, nn.CrossEntropyLoss on the softmax
    OUTPUT, as the reference does).  Returns losses, outputs, and the gradients of every decoder parameter and of the
    [SEG] embeddings."""
    names = [k for k in w if k.startswith("mask_decoder_")]
    wv = {k: (v.detach().clone().requires_grad_(True) if k in names else v.detach()) for k, v in w.items()}
    segs = [s.detach().clone().requires_grad_(True) for s in seg_embeds]
    pl, pr, taxes, low_l, low_r = [], [], [], [], []
    for i, s in enumerate(segs):
        dl = O.decode(wv, "left", emb[i], s)
        dr = O.decode(wv, "right", emb[i], s)
        low_l.append(dl["masks"])
        low_r.append(dr["masks"])
        taxes.append(dl["taxonomy"])
        pl.append(O.postprocess_masks(dl["masks"], resize_list[i], original_list[i])[:, 0])
        pr.append(O.postprocess_masks(dr["masks"], resize_list[i], original_list[i])[:, 0])
    losses = sam_losses(pl, pr, gt_left, gt_right, gt_taxonomies)
    tax = 0.0
    for i, t in enumerate(taxes):     # LISA.py:411-415: CrossEntropyLoss(pred_taxonomy (N_i, 4), gt (1, 4)) needs N_i == 1
        tax = tax + torch.nn.functional.cross_entropy(t, gt_taxonomies[i][None].float().expand(t.shape[0], -1))
    tax = tax / len(taxes)
    loss = losses["mask_loss"] + tax_weight * tax
    loss.backward()
    grads = {k: wv[k].grad for k in names}
    return {"loss": loss.detach(), "taxonomy_ce_loss": tax.detach(),
            **{k: v.detach() for k, v in losses.items()},
            "low_res_left": torch.cat(low_l).detach(), "low_res_right": torch.cat(low_r).detach(),
            "taxonomy": torch.cat(taxes).detach(), "grads": grads, "d_seg": [s.grad for s in segs]}


def grad_digest(g: torch.Tensor, k: int = 16) -> torch.Tensor:
    """[sum, L2 norm, max |.|, k values at fixed pseudo-random positions] of a gradient tensor (what the golden file
    keeps per parameter; the decoders have 4 M parameters each)."""
    f = g.detach().double().flatten()
    n = f.numel()
    idx = (torch.arange(k, dtype=torch.int64) * 2654435761 + 12345) % n
    return torch.cat([torch.stack([f.sum(), f.norm(), f.abs().max()]), f[idx]]).float()
