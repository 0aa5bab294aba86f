"""The callers either side of the SAM branch, on the device.  Mixed into `B200Sam`.

* `preprocess_frames`  replaces `ResizeLongestSide.apply_image` + `preprocess` + the dtype cast of the reference's
  callers (2Haff/inference.py:244-256, chat.py:193-205; segment_anything/utils/transforms.py:27-34) - raw uint8 RGB
  frames in, normalised zero-padded 1024^2 tensors out, bit-exact against PIL + torch.
* `seg_project`        replaces `text_hidden_fcs[0](hidden)[seg_token_mask]` + the offset bookkeeping of
  `LISAForCausalLM.evaluate` (2Haff/model/LISA.py:457-485), computing the projection on the `[SEG]` rows only.  It takes
  the reference's own `model.text_hidden_fcs` ModuleList (its parameters are read in place), so nothing about the
  LLM side of the checkpoint changes.
* `mask_metrics`       replaces the host-side numpy evaluation of `train_ds.py::validate` (train_ds.py:662-709,
  761-796): taxonomy gate, IoU and IoCM from pixel counts reduced on the device.

All arithmetic is in libb200sam; there is no PyTorch / CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib

OUT_CH = 256
_DT = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16}


def _sp() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def get_preprocess_shape(oldh: int, oldw: int, long_side_length: int = 1024) -> Tuple[int, int]:
    """ResizeLongestSide.get_preprocess_shape (transforms.py:103-113) - host arithmetic only."""
    scale = long_side_length * 1.0 / max(oldh, oldw)
    return int(oldh * scale + 0.5), int(oldw * scale + 0.5)


class CallersMixin:
    # ------------------------------------------------------------------ row 1: frames
    @torch.no_grad()
    def preprocess_frames(self, frames_rgb: torch.Tensor, dtype: torch.dtype = torch.bfloat16
                          ) -> Tuple[torch.Tensor, Tuple[int, int]]:
        """frames_rgb: (B,H,W,3) or (H,W,3) uint8 RGB on the GPU (what cv2.imread + cvtColor yields,
        inference.py:229-230).  Returns (`images` (B,3,1024,1024) in `dtype`, `resize` (h,w)): exactly the tensor and the
        `resize_list` entry the reference feeds to `evaluate` (inference.py:244-256)."""
        if frames_rgb.dim() == 3:
            frames_rgb = frames_rgb[None]
        if frames_rgb.dim() != 4 or frames_rgb.shape[-1] != 3 or frames_rgb.dtype != torch.uint8:
            raise ValueError(f"frames must be (B,H,W,3) uint8 RGB, got {tuple(frames_rgb.shape)} {frames_rgb.dtype}")
        if dtype not in _DT:
            raise TypeError(f"unsupported output dtype {dtype}")
        dev = frames_rgb.device
        h = self._ensure_handle_only(dev)
        B, H, W, _ = frames_rgb.shape
        frames_rgb = frames_rgb.contiguous()
        out = torch.empty((B, 3, 1024, 1024), dtype=dtype, device=dev)
        hw = (C.c_int * 2)()
        with torch.cuda.device(dev):
            rc = _lib.load().b200sam_preprocess(h, _p(frames_rgb), B, H, W, _p(out), _DT[dtype], hw, _sp())
        _lib.check(rc, h, "preprocess")
        return out, (int(hw[0]), int(hw[1]))

    # ------------------------------------------------------------------ row 2: [SEG] rows
    @torch.no_grad()
    def seg_project(self, hidden_states: torch.Tensor, output_ids: torch.Tensor, seg_token_idx: int,
                    text_hidden_fcs, max_prompts: Optional[int] = None) -> Tuple[List[torch.Tensor], List[int]]:
        """LISA.py:457-485.  hidden_states (b,L,h) = `outputs.hidden_states[-1]`, output_ids (b,T) =
        `outputs.sequences`, text_hidden_fcs = the reference's `model.text_hidden_fcs` (ModuleList of one
        Sequential(Linear, ReLU, Linear, Dropout)).  Returns (`pred_embeddings`: list of b tensors (N_i,256) fp32,
        offsets (b+1)).  One (b+1)-int device->host read, as the reference's `.cumsum()` indexing implies."""
        if len(text_hidden_fcs) != 1:                     # LISA.py:470
            raise AssertionError("len(text_hidden_fcs) must be 1")
        fc1, fc2 = text_hidden_fcs[0][0], text_hidden_fcs[0][2]
        if hidden_states.dim() != 3 or output_ids.dim() != 2 or hidden_states.shape[0] != output_ids.shape[0]:
            raise ValueError("hidden_states must be (b,L,h) and output_ids (b,T)")
        b, L, hdim = hidden_states.shape
        T = output_ids.shape[1]
        if tuple(fc1.weight.shape) != (hdim, hdim) or tuple(fc2.weight.shape) != (OUT_CH, hdim):
            raise ValueError("text_hidden_fcs does not match the hidden size")
        if hidden_states.dtype not in _DT:
            raise TypeError(f"unsupported hidden dtype {hidden_states.dtype}")
        dev = hidden_states.device
        h = self._ensure_handle_only(dev)
        wdt = fc1.weight.dtype
        if wdt == torch.float16:
            wdt = torch.float32
        ws = [t.detach().to(device=dev, dtype=wdt).contiguous() for t in (fc1.weight, fc1.bias, fc2.weight, fc2.bias)]
        cap = int(max_prompts) if max_prompts else max(self._cap[1], 32)
        hidden_states = hidden_states.contiguous()
        ids = output_ids.to(device=dev, dtype=torch.int64).contiguous()
        emb = torch.empty((cap, OUT_CH), dtype=torch.float32, device=dev)
        counts = torch.empty((b + 1,), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.load().b200sam_seg_project(h, _p(hidden_states), _DT[hidden_states.dtype], b, L, hdim, _p(ids), T,
                                                 int(seg_token_idx), _p(ws[0]), _p(ws[1]), _p(ws[2]), _p(ws[3]),
                                                 _DT[wdt], _p(emb), cap, _p(counts), _sp())
        _lib.check(rc, h, "seg_project")
        c = counts.tolist()
        if c[b] > cap:
            raise RuntimeError(f"{c[b]} [SEG] tokens exceed max_prompts={cap}")
        offsets = [0]
        for n in c[:b]:
            offsets.append(offsets[-1] + n)
        return [emb[offsets[i]:offsets[i + 1]] for i in range(b)], offsets

    # ------------------------------------------------------------------ row 3: evaluation
    @torch.no_grad()
    def mask_metrics(self, masks_left: torch.Tensor, masks_right: torch.Tensor, taxonomy: torch.Tensor,
                     gt: torch.Tensor, gt_index: Optional[Sequence[int]] = None, return_pred: bool = False
                     ) -> Dict[str, torch.Tensor]:
        """train_ds.py:662-709.  masks_left / masks_right (P,H,W) uint8 thresholded masks, taxonomy (P,4), gt (G,H,W)
        uint8 (`gt_aff` = OR of the two ground-truth masks, train_ds.py:659), prompt p is scored against
        gt[gt_index[p]] (default gt[p]).  Returns `counts` (P,3) int64 {intersection, union, area}, `iou` and `iocm`
        (P,) float64 with the reference's formulas - including its `comparison_mask.sum()` over a {0,255} mask
        (train_ds.py:792), i.e. IoCM = I / (255 * area) - and `pred` (P,H,W) uint8 {0,255} if asked."""
        if masks_left.dtype != torch.uint8 or masks_right.dtype != torch.uint8 or gt.dtype != torch.uint8:
            raise TypeError("masks and gt must be uint8")
        if masks_left.dim() != 3 or masks_left.shape != masks_right.shape or gt.dim() != 3 or \
                gt.shape[1:] != masks_left.shape[1:]:
            raise ValueError("masks must be (P,H,W) and gt (G,H,W) of one size")
        P, H, W = masks_left.shape
        if tuple(taxonomy.shape) != (P, 4):
            raise ValueError("taxonomy must be (P,4)")
        dev = masks_left.device
        h = self._ensure_handle_only(dev)
        idx = None
        if gt_index is not None:
            if len(gt_index) != P or any(not (0 <= int(i) < gt.shape[0]) for i in gt_index):
                raise ValueError("gt_index must hold one valid gt row per prompt")
            idx = torch.tensor([int(i) for i in gt_index], dtype=torch.int32, device=dev)
        elif gt.shape[0] != P:
            raise ValueError("gt must have one mask per prompt when gt_index is not given")
        counts = torch.empty((P, 3), dtype=torch.int64, device=dev)
        pred = torch.empty((P, H, W), dtype=torch.uint8, device=dev) if return_pred else None
        tax = taxonomy.to(device=dev, dtype=torch.float32).contiguous()
        with torch.cuda.device(dev):
            rc = _lib.load().b200sam_mask_metrics(h, _p(masks_left.contiguous()), _p(masks_right.contiguous()), _p(tax),
                                                  _p(gt.contiguous()), _p(idx), P, H, W, _p(counts), _p(pred), _sp())
        _lib.check(rc, h, "mask_metrics")
        inter, union, area = (counts[:, i].to(torch.float64) for i in range(3))
        zero = torch.zeros_like(inter)
        out = {"counts": counts,
               "iou": torch.where(union != 0, inter / union.clamp(min=1), zero),                 # train_ds.py:775-776
               "iocm": torch.where(area != 0, inter / (255.0 * area).clamp(min=1), zero)}        # train_ds.py:792-794
        if return_pred:
            out["pred"] = pred
        return out
