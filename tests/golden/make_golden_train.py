"""Golden fixture for the training forward / backward of the SAM branch, from the UNMODIFIED reference.

Run in the build container only (/root/reference does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_train.py

The reference's own modules (segment_anything MaskDecoder x 2, PromptEncoder, Sam.postprocess_masks) run under autograd
on the CPU in fp32 with the seeded weights of oracle/synth.py, in train() mode with requires_grad decoder parameters
exactly as LisaMetaModel.initialize_lisa_modules sets them up (2Haff/model/LISA.py:79-90), following the per-image loop
of LISA.model_forward (LISA.py:284-325).  dice_loss and sigmoid_ce_loss are the reference's own functions: LISA.py cannot
be imported here (it pulls in the LLaVA model code), so the two function definitions are compiled from the reference
file's source at run time - nothing is copied into the repo.  The few lines that combine them (taxonomy gate,
per-image accumulation, division by num_masks; LISA.py:346-420) live inside LisaForCausalLM.model_forward behind the
LLM call and are restated below.

Stored in tests/golden/train_tiny.npz: the losses, the low-res logits and taxonomy outputs, the full gradient of the
[SEG] embeddings, and a digest (sum, norm, max, 16 samples) of the gradient of every decoder parameter.
"""
from __future__ import annotations

import ast
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/2Haff/model")

from oracle import synth  # noqa: E402
from oracle import train_oracle as TO  # noqa: E402
from tests.golden.make_golden import build_reference  # noqa: E402

LISA_PY = "/root/reference/2Haff/model/LISA.py"


def reference_loss_functions():
    tree = ast.parse(open(LISA_PY).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("dice_loss", "sigmoid_ce_loss")]
    assert len(keep) == 2
    ns = {"torch": torch, "F": torch.nn.functional}
    exec(compile(ast.Module(body=keep, type_ignores=[]), LISA_PY, "exec"), ns)
    return ns["dice_loss"], ns["sigmoid_ce_loss"]


def main():
    torch.set_num_threads(os.cpu_count())
    dice_loss, sigmoid_ce_loss = reference_loss_functions()
    sam = build_reference(synth.ARCHS["tiny64"])
    # LISA.py:81-90
    for p in sam.parameters():
        p.requires_grad = False
    for dec in (sam.mask_decoder_left, sam.mask_decoder_right):
        dec.train()
        for p in dec.parameters():
            p.requires_grad = True
    emb, segs, gt_left, gt_right, tax = synth.make_train_inputs()
    segs = [s.clone().requires_grad_(True) for s in segs]
    case = synth.TRAIN_CASE
    pl, pr, taxes, low_l, low_r = [], [], [], [], []
    for i, s in enumerate(segs):                                  # LISA.py:284-325
        sparse, dense = sam.prompt_encoder(points=None, boxes=None, masks=None, text_embeds=s.unsqueeze(1))
        sparse = sparse.to(s.dtype)
        ml, _, t = sam.mask_decoder_left(image_embeddings=emb[i].unsqueeze(0), image_pe=sam.prompt_encoder.get_dense_pe(),
                                         sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense,
                                         multimask_output=False)
        mr, _ = sam.mask_decoder_right(image_embeddings=emb[i].unsqueeze(0), image_pe=sam.prompt_encoder.get_dense_pe(),
                                       sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense,
                                       multimask_output=False)
        taxes.append(t)
        low_l.append(ml)
        low_r.append(mr)
        pl.append(sam.postprocess_masks(ml, input_size=case["resize"][i], original_size=case["original"][i])[:, 0])
        pr.append(sam.postprocess_masks(mr, input_size=case["resize"][i], original_size=case["original"][i])[:, 0])
    # LISA.py:346-420 restated (see the module docstring); bce_loss_weight 2.0 / dice_loss_weight 0.5 = train_ds.py defaults
    wl = tax[:, 0] + tax[:, 2] + tax[:, 3]
    wr = tax[:, 1] + tax[:, 2] + tax[:, 3]
    bl = br = dl = dr = 0
    n = 0
    ce = torch.nn.CrossEntropyLoss(reduction="mean")            # LISA.py:151
    tce = 0
    for i in range(len(pl)):
        a, b = wl[i] * pl[i], wr[i] * pr[i]
        k = gt_left[i].shape[0]
        bl = bl + sigmoid_ce_loss(a, gt_left[i], num_masks=k) * k
        dl = dl + dice_loss(a, gt_left[i], num_masks=k) * k
        br = br + sigmoid_ce_loss(b, gt_right[i], num_masks=k) * k
        dr = dr + dice_loss(b, gt_right[i], num_masks=k) * k
        n += k
        tce = tce + ce(taxes[i], tax[i][None].float().expand(taxes[i].shape[0], -1))
    tce = tce / len(pl)
    bce = 2.0 * bl / (n + 1e-8) + 2.0 * br / (n + 1e-8)
    dice = 0.5 * dl / (n + 1e-8) + 0.5 * dr / (n + 1e-8)
    loss = bce + dice + tce
    loss.backward()
    out = {"loss": loss.detach().numpy(), "mask_bce_loss": bce.detach().numpy(), "mask_dice_loss": dice.detach().numpy(),
           "taxonomy_ce_loss": tce.detach().numpy(),
           "low_res_left": torch.cat(low_l).detach().numpy(), "low_res_right": torch.cat(low_r).detach().numpy(),
           "taxonomy": torch.cat(taxes).detach().numpy()}
    for i, s in enumerate(segs):
        out[f"d_seg{i}"] = s.grad.numpy()
    names = []
    for side, dec in (("left", sam.mask_decoder_left), ("right", sam.mask_decoder_right)):
        for k, p in dec.named_parameters():
            name = f"mask_decoder_{side}.{k}"
            names.append(name)
            if p.grad is None:       # the IoU head takes no part in the loss (LISA.py never uses iou_predictions)
                assert "iou_prediction_head" in k or k == "iou_token.weight", k
                out["g:" + name] = np.zeros((0,), np.float32)
            else:
                out["g:" + name] = TO.grad_digest(p.grad).numpy()
    out["param_names"] = np.array(names)
    path = os.path.join(ROOT, "tests", "golden", "train_tiny.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; loss", float(loss), "bce", float(bce), "dice", float(dice),
          "tax", float(tce))


if __name__ == "__main__":
    main()
