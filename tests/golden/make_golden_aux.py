"""Golden fixtures for the rows either side of the SAM branch, made by running the
UNMODIFIED reference in the build container (never on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_aux.py

* preprocess_<HxW>.npz - `ResizeLongestSide.apply_image` (transforms.py:27-34) + `preprocess` (inference.py:91-105,
  exec'd from the source text because inference.py imports deepspeed-era modules): SHA-256 of the resized uint8 image
  and of the padded fp32 tensor, plus 4096 sampled values (full tensors are 12 MB each; the hash pins every bit).
* aux_segproj.npz - the reference's `text_hidden_fcs` construction (LISA.py:92-101) and the `[SEG]` mask / gather /
  offset statements of `evaluate` (LISA.py:457-485), both exec'd from the source text of LISA.py with a stub `self`
  (`from model.LISA import ...` fails under transformers 5) and `.cuda()` made a no-op.
* aux_metrics.npz - `calculate_iou` / `calculate_iocm` (train_ds.py:761-796) and the taxonomy gate
  (train_ds.py:690-706) exec'd from the source text of train_ds.py.

Nothing of the reference is copied into the repo: only its numeric outputs.
"""
from __future__ import annotations

import hashlib
import os
import sys
import textwrap
import types

import numpy as np
import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/2Haff"
sys.path.insert(0, REF + "/model")

from oracle.synth import synth_image  # noqa: E402

PRE_CASES = [(1080, 1920, 0), (480, 640, 1), (1024, 1024, 2), (300, 200, 3), (720, 1280, 4), (37, 53, 5),
             (1500, 1024, 6), (2160, 3840, 7)]


def src_lines(path, lo, hi):
    with open(path) as fh:
        lines = fh.readlines()
    return textwrap.dedent("".join(lines[lo - 1:hi]))


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def make_preprocess():
    from segment_anything.utils.transforms import ResizeLongestSide
    ns = {"torch": torch, "F": torch.nn.functional}
    exec(src_lines(REF + "/inference.py", 91, 105), ns)          # def preprocess(...)
    transform = ResizeLongestSide(1024)
    for H, W, seed in PRE_CASES:
        img = synth_image(H, W, seed)
        small = transform.apply_image(img)
        x = ns["preprocess"](torch.from_numpy(small).permute(2, 0, 1).contiguous()).numpy()
        assert x.shape == (3, 1024, 1024) and x.dtype == np.float32
        rng = np.random.default_rng(100 + seed)
        idx = rng.integers(0, x.size, 4096)
        np.savez_compressed(os.path.join(HERE, f"preprocess_{H}x{W}.npz"), hw=np.array([H, W]), seed=np.array(seed),
                            resize=np.array(small.shape[:2]), sha_u8=np.array(sha(small)), sha_f32=np.array(sha(x)),
                            sha_bf16=np.array(sha(torch.from_numpy(x).bfloat16().view(torch.int16).numpy())),
                            sample_idx=idx, sample_val=x.reshape(-1)[idx])
        print("preprocess", (H, W), "->", small.shape[:2], sha(x)[:12])


def make_segproj():
    h, out_dim, seg_idx = 512, 256, 32003
    torch.manual_seed(7)
    cfg = types.SimpleNamespace(hidden_size=h, out_dim=out_dim)
    me = nn.Module()
    exec(src_lines(REF + "/model/LISA.py", 92, 101), {"nn": nn, "config": cfg, "self": me})
    fcs = me.text_hidden_fcs.eval()
    b, T = 3, 40
    g = torch.Generator().manual_seed(8)
    output_ids = torch.randint(0, 32000, (b, T), generator=g)
    output_ids[0, [5, 17, 30]] = seg_idx       # three prompts
    output_ids[0, 0] = seg_idx                 # position 0 is dropped by [:, 1:]
    output_ids[2, [T - 1]] = seg_idx           # last token; row 1 has no [SEG] at all
    hidden = torch.randn((b, 255 + T - 1, h), generator=g)
    model_stub = types.SimpleNamespace(text_hidden_fcs=fcs)
    self_stub = types.SimpleNamespace(seg_token_idx=seg_idx, model=model_stub)
    ns = {"torch": torch, "self": self_stub, "output_ids": output_ids, "output_hidden_states": hidden}
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        with torch.no_grad():
            exec(src_lines(REF + "/model/LISA.py", 457, 485), ns)
    finally:
        torch.Tensor.cuda = orig_cuda
    pred = ns["pred_embeddings"]
    sd = {k: v.detach().numpy() for k, v in fcs[0].state_dict().items()}
    np.savez_compressed(os.path.join(HERE, "aux_segproj.npz"), seg_idx=np.array(seg_idx), output_ids=output_ids.numpy(),
                        hidden=hidden.numpy(), offsets=ns["seg_token_offset"].numpy(),
                        pred=torch.cat(pred, 0).numpy(), **{"w_" + k.replace(".", "_"): v for k, v in sd.items()})
    print("segproj", [tuple(p.shape) for p in pred], ns["seg_token_offset"].tolist())


def make_metrics():
    import cv2  # noqa: F401  (the gate below uses cv2.bitwise_or like train_ds.py:704)
    ns = {"np": np}
    exec(src_lines(REF + "/train_ds.py", 761, 796), ns)
    gate_src = src_lines(REF + "/train_ds.py", 690, 706)
    H, W = 135, 240
    rng = np.random.default_rng(11)
    yy, xx = np.mgrid[0:H, 0:W]
    rec = {k: [] for k in ("left", "right", "gt", "tax", "pred", "iou", "iocm")}
    for case in range(8):
        cx, cy, r = rng.integers(40, 200), rng.integers(30, 100), rng.integers(10, 50)
        left = ((xx - cx) ** 2 + (yy - cy) ** 2 < r * r).astype(np.int32)
        right = ((xx - cx - 30) ** 2 + (yy - cy + 10) ** 2 < (r + 5) ** 2).astype(np.int32)
        gt = ((xx - cx - 10) ** 2 + (yy - cy) ** 2 < (r + 3) ** 2).astype(np.int32)
        if case == 6:
            left[:] = 0
            right[:] = 0                      # empty prediction: both ratios fall back to 0.0
        if case == 7:
            gt[:] = 0
        tax = np.zeros(4, np.float32)
        tax[case % 4] = 1.0
        tax += rng.random(4).astype(np.float32) * 0.1
        g = {"torch": torch, "cv2": cv2, "taxonomy": torch.from_numpy(tax),
             "pred_masks_left": torch.from_numpy(left)[None], "pred_masks_right": torch.from_numpy(right)[None]}
        exec(gate_src, g)
        pred = g["pred_mask"]
        rec["left"].append(left.astype(np.uint8)), rec["right"].append(right.astype(np.uint8))
        rec["gt"].append(gt.astype(np.uint8)), rec["tax"].append(tax), rec["pred"].append(pred.astype(np.int32))
        rec["iou"].append(float(ns["calculate_iou"](pred, gt)))
        rec["iocm"].append(float(ns["calculate_iocm"](gt, pred)))
        print("metrics", case, "argmax", int(tax.argmax()), rec["iou"][-1], rec["iocm"][-1])
    np.savez_compressed(os.path.join(HERE, "aux_metrics.npz"), **{k: np.stack(v) for k, v in rec.items()})


if __name__ == "__main__":
    which = sys.argv[1:] or ["preprocess", "segproj", "metrics"]
    for w in which:
        {"preprocess": make_preprocess, "segproj": make_segproj, "metrics": make_metrics}[w]()
