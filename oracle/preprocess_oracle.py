"""TEST INFRASTRUCTURE ONLY - CPU restatement of the frame preprocessing of 2HandedAfforder.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product path
(2handedafforder_b200/) never does.

Path restated (all integer / exactly-rounded fp32 arithmetic, so the CUDA kernel is compared BIT-EXACTLY):

  1. `ResizeLongestSide.get_preprocess_shape`  2Haff/model/segment_anything/utils/transforms.py:103-113
  2. `ResizeLongestSide.apply_image`           transforms.py:27-34 = torchvision `resize(to_pil_image(img), size)` =
     PIL `Image.resize(size, BILINEAR)` on an RGB uint8 image.  The arithmetic lives in a third-party dependency that
     is not vendored in the reference: Pillow (pinned `pillow==9.4.0` by 2Haff/requirements.txt's torchvision
     0.14.1 stack; this container has Pillow 12.2.0 - the 8-bit resample algorithm, `src/libImaging/Resample.c`,
     is unchanged between them).  Published algorithm, restated below from its description:
       * two separable passes, horizontal first, each rounding to uint8;
       * the horizontal pass only touches the input rows the vertical pass needs;
       * per output index: centre = (i + 0.5) * scale, support = max(scale, 1), window
         [int(centre - support + 0.5), int(centre + support + 0.5)) clipped to the image, triangle weights
         normalised to sum 1 in double precision, then converted to 22-bit fixed point with round-half-away;
       * accumulate in int32 starting from 1 << 21, shift right by 22, clip to [0, 255].
  3. `preprocess`  2Haff/inference.py:91-105 (= chat.py:49-63, sam.py:191-201): (x - mean) / std in fp32 on the CHW
     tensor, zero pad bottom/right to 1024x1024; callers then cast to the model dtype (inference.py:251-256).

Pinned by tests/golden/preprocess_*.npz: outputs of the reference's own `ResizeLongestSide.apply_image` +
`preprocess` (imported from /root/reference in the build container by tests/golden/make_golden_aux.py).
"""
from __future__ import annotations

import math
from typing import Tuple

import numpy as np

PRECISION_BITS = 32 - 8 - 2
PIXEL_MEAN = np.array([123.675, 116.28, 103.53], dtype=np.float32)   # inference.py:93
PIXEL_STD = np.array([58.395, 57.12, 57.375], dtype=np.float32)      # inference.py:94
IMG_SIZE = 1024


def get_preprocess_shape(oldh: int, oldw: int, long_side: int = IMG_SIZE) -> Tuple[int, int]:
    """transforms.py:103-113."""
    scale = long_side * 1.0 / max(oldh, oldw)
    newh, neww = oldh * scale, oldw * scale
    return int(newh + 0.5), int(neww + 0.5)


def resample_coeffs(in_size: int, out_size: int):
    """Window start, window length and 22-bit fixed-point triangle weights per output index (Pillow
    precompute_coeffs + normalize_coeffs_8bpc, bilinear filter, box = the whole image)."""
    scale = float(in_size) / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    xmin = np.zeros(out_size, dtype=np.int32)
    xcnt = np.zeros(out_size, dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        lo = int(center - support + 0.5)
        lo = max(lo, 0)
        hi = int(center + support + 0.5)
        hi = min(hi, in_size)
        n = hi - lo
        w = np.zeros(ksize, dtype=np.float64)
        for x in range(n):
            t = abs((x + lo - center + 0.5) * ss)
            w[x] = 1.0 - t if t < 1.0 else 0.0
        tot = 0.0                                  # Pillow accumulates ww sequentially in index order
        for x in range(n):
            tot += w[x]
        if tot != 0.0:
            for x in range(n):
                w[x] /= tot
        for x in range(ksize):
            v = w[x] * (1 << PRECISION_BITS)
            kk[xx, x] = int(-0.5 + v) if w[x] < 0 else int(0.5 + v)
        xmin[xx], xcnt[xx] = lo, n
    return xmin, xcnt, kk


def _pass(img: np.ndarray, axis_out: int, xmin, xcnt, kk) -> np.ndarray:
    """One separable pass along axis 1 of img (rows, in, C) uint8 -> (rows, out, C) uint8."""
    rows, _, ch = img.shape
    out = np.empty((rows, axis_out, ch), dtype=np.uint8)
    src = img.astype(np.int64)
    for xx in range(axis_out):
        n, lo = int(xcnt[xx]), int(xmin[xx])
        acc = np.full((rows, ch), 1 << (PRECISION_BITS - 1), dtype=np.int64)
        for x in range(n):
            acc += src[:, lo + x, :] * int(kk[xx, x])
        # int32 wrap cannot happen (weights sum to 2^22, pixels <= 255); arithmetic shift, then clip8
        out[:, xx, :] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return out


def pil_resize_bilinear(img: np.ndarray, new_h: int, new_w: int) -> np.ndarray:
    """(H,W,3) uint8 -> (new_h,new_w,3) uint8, Pillow ImagingResample with the bilinear filter."""
    H, W, _ = img.shape
    need_h, need_v = new_w != W, new_h != H
    out = img
    if need_v:
        ymin, ycnt, ky = resample_coeffs(H, new_h)
    if need_h:
        xmin, xcnt, kx = resample_coeffs(W, new_w)
        first, last = (int(ymin[0]), int(ymin[-1] + ycnt[-1])) if need_v else (0, H)
        out = _pass(np.ascontiguousarray(out[first:last]), new_w, xmin, xcnt, kx)
        if need_v:
            ymin = ymin - first
    if need_v:
        out = _pass(np.ascontiguousarray(out.transpose(1, 0, 2)), new_h, ymin, ycnt, ky).transpose(1, 0, 2)
    return np.ascontiguousarray(out)


def preprocess_frame(img: np.ndarray, img_size: int = IMG_SIZE):
    """(H,W,3) uint8 RGB -> ((3,img_size,img_size) float32 normalised + zero padded, (new_h,new_w)).
    transforms.py:27-34 then inference.py:91-105."""
    new_h, new_w = get_preprocess_shape(img.shape[0], img.shape[1], img_size)
    small = pil_resize_bilinear(img, new_h, new_w)
    x = small.transpose(2, 0, 1).astype(np.float32)
    x = (x - PIXEL_MEAN[:, None, None]) / PIXEL_STD[:, None, None]      # fp32 subtract, fp32 IEEE divide
    out = np.zeros((3, img_size, img_size), dtype=np.float32)
    out[:, :new_h, :new_w] = x
    return out, (new_h, new_w)
