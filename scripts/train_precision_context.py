#!/usr/bin/env python
"""How far the reference's OWN training precision is from fp32 on the decoder gradients (CPU, build container or box):
oracle/train_oracle.py's step on tests/golden's training case in fp32 and under torch.autocast(bfloat16) - train_ds.py
trains with --precision bf16.  Context for the tolerance of the TF32 tensor-core path (tests/test_gpu_train.py).
Measurement aid; writes profiles/r4_train_precision_context.json."""
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import synth, train_oracle as TO  # noqa: E402

w = synth.make_state_dict(synth.ARCHS["tiny64"], seed=0)
emb, segs, gl, gr, tax = synth.make_train_inputs()
c = synth.TRAIN_CASE
r = TO.train_step(w, emb, segs, c["resize"], c["original"], gl, gr, tax)
with torch.autocast("cpu", dtype=torch.bfloat16):
    rb = TO.train_step(w, emb, segs, c["resize"], c["original"], gl, gr, tax)
gmax = max(float(v.abs().max()) for v in r["grads"].values() if v is not None)
rows = []
for k, v in r["grads"].items():
    if v is None or float(v.abs().max()) < 1e-6 * gmax:      # shift-invariant biases: exact zeros up to rounding
        continue
    e, m = float((rb["grads"][k].float() - v).abs().max()), float(v.abs().max())
    rows.append((e / m, e / gmax, k))
rows.sort(reverse=True)
rep = {"what": "bf16 autocast (the reference's training precision) vs fp32, same CPU oracle, tests/golden training case",
       "tensors": len(rows), "largest_gradient": gmax,
       "param_grads_max_rel_err_own_max": rows[0][0], "worst_tensor": rows[0][2],
       "param_grads_median_rel_err": statistics.median(x[0] for x in rows),
       "param_grads_max_err_over_largest_gradient": max(x[1] for x in rows),
       "d_seg_rel_err": [float((rb["d_seg"][i].float() - d).abs().max() / d.abs().max()) for i, d in enumerate(r["d_seg"])],
       "loss_fp32": float(r["loss"]), "loss_bf16": float(rb["loss"])}
print(json.dumps(rep, indent=1))
with open(os.path.join(ROOT, "profiles", "r4_train_precision_context.json"), "w") as fh:
    json.dump(rep, fh, indent=1)
