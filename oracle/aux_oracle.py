"""TEST INFRASTRUCTURE ONLY - CPU restatement of the two callers either side of the SAM branch.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product path never does.

Row 2 - `[SEG]` projection and gather, 2Haff/model/LISA.py:
  * `text_hidden_fcs[0]` = Sequential(Linear(h,h), ReLU, Linear(h,256), Dropout(0))        LISA.py:95-101
  * evaluate(): seg mask = (output_ids[:,1:] == seg_token_idx) left-padded by 255 False columns (one image token
    expands to 256 CLIP patches), projection of ALL rows, boolean gather, per-image offsets  LISA.py:457-485
Row 3 - evaluation metrics of train_ds.py::validate:
  * threshold `> 0`, taxonomy argmax gate (0 -> left, 1 -> right, else left OR right), masks as {0,255}
    train_ds.py:662-663,690-706
  * `calculate_iou`  train_ds.py:761-777 ; `calculate_iocm`  train_ds.py:779-796.  NOTE the reference divides the
    intersection COUNT by `comparison_mask.sum()` where the predicted mask holds 0/255, i.e. IoCM = I / (255 * |pred|):
    restated as is (a drop-in must print the numbers the reference prints).

Pinned by tests/golden/aux_*.npz: outputs of the reference's own modules / functions on seeded inputs
(tests/golden/make_golden_aux.py; train_ds.py cannot be imported without deepspeed, so its two metric functions are
exec'd from the source text of /root/reference/2Haff/train_ds.py there).
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

IMAGE_TOKEN_PAD = 255          # LISA.py:461-466


def seg_rows(output_ids: np.ndarray, seg_token_idx: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """LISA.py:457-476.  output_ids (b,T) int -> (batch index, row index into the (b, 255+T-1, h) hidden states) of
    every [SEG] token in row-major order, and offsets (b+1,)."""
    mask = output_ids[:, 1:] == seg_token_idx
    mask = np.concatenate([np.zeros((mask.shape[0], IMAGE_TOKEN_PAD), dtype=bool), mask], axis=1)
    bi, ri = np.nonzero(mask)
    counts = mask.sum(-1)
    offsets = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    return bi, ri, offsets


def text_hidden_fc(x: np.ndarray, w1, b1, w2, b2) -> np.ndarray:
    """LISA.py:95-101 on rows x (n,h), fp32."""
    x = x.astype(np.float32)
    y = np.maximum(x @ w1.astype(np.float32).T + b1.astype(np.float32), 0.0)
    return y @ w2.astype(np.float32).T + b2.astype(np.float32)


def seg_project(hidden: np.ndarray, output_ids: np.ndarray, seg_token_idx: int, w1, b1, w2, b2):
    """LISA.py:457-485: (b,L,h) last hidden states -> list of (N_i,256) prompt embeddings (projection of every row,
    then the boolean gather; gathering first gives the same rows)."""
    bi, ri, offsets = seg_rows(output_ids, seg_token_idx)
    rows = hidden[bi, ri] if len(bi) else np.zeros((0, hidden.shape[-1]), np.float32)
    emb = text_hidden_fc(rows, w1, b1, w2, b2)
    return [emb[offsets[i]:offsets[i + 1]] for i in range(len(offsets) - 1)], offsets


def gated_mask(mask_left: np.ndarray, mask_right: np.ndarray, taxonomy: np.ndarray) -> np.ndarray:
    """train_ds.py:690-706: u8 {0,1} masks (H,W) + taxonomy (4,) -> int32 {0,255} mask."""
    k = int(np.argmax(taxonomy))
    left = np.where(mask_left > 0, 255, 0).astype(np.int32)
    right = np.where(mask_right > 0, 255, 0).astype(np.int32)
    if k == 0:
        return left
    if k == 1:
        return right
    return left | right


def mask_counts(pred255: np.ndarray, gt: np.ndarray) -> Tuple[int, int, int]:
    """(intersection, union, predicted area) pixel counts of train_ds.py:761-796."""
    p, g = pred255 != 0, gt != 0
    return int((p & g).sum()), int((p | g).sum()), int(p.sum())


def iou_iocm(pred255: np.ndarray, gt: np.ndarray) -> Tuple[float, float]:
    """calculate_iou(pred, gt) and calculate_iocm(gt, pred) exactly as train_ds.py:708-709 calls them."""
    inter, union, area = mask_counts(pred255, gt)
    iou = inter / union if union != 0 else 0.0
    comparison_area = int(pred255.astype(np.int64).sum())          # = 255 * area (reference quirk, see header)
    iocm = inter / comparison_area if comparison_area != 0 else 0.0
    return iou, iocm


# ---------------------------------------------------------------------------------------------- wire format

def pack_bits(mask: np.ndarray) -> np.ndarray:
    """(..., H, W) {0,1} -> (..., H, ceil(W/8)) u8: the 1-bit wire format of b200sam_postprocess_bits (bit 7 of byte
    X/8 = pixel 8*(X/8), pad bits 0) - numpy.packbits along the row, which is also how the golden fixtures store masks."""
    return np.packbits(mask.astype(bool), axis=-1)


_MASK64 = (1 << 64) - 1


def _mix64(z: int) -> int:
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK64
    return z ^ (z >> 31)


def checksum64(data: np.ndarray) -> np.ndarray:
    """include/b200sam.h b200sam_checksum64: per item of a (n, ...) u8 array, the sum over its 8-byte little-endian
    words w_k (tail zero padded) of splitmix64_final(w_k ^ (k+1) * 0x9E3779B97F4A7C15) mod 2^64, returned as int64 bit
    patterns.  Vectorised with numpy uint64 wrap-around arithmetic."""
    d = np.ascontiguousarray(data).reshape(data.shape[0], -1).astype(np.uint8)
    n, nb = d.shape
    pad = (-nb) % 8
    if pad:
        d = np.concatenate([d, np.zeros((n, pad), np.uint8)], axis=1)
    w = d.view("<u8")
    k = (np.arange(1, w.shape[1] + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
    with np.errstate(over="ignore"):
        z = w ^ k
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        s = z.sum(axis=1, dtype=np.uint64)
    return s.view(np.int64)
