"""CPU checks that pin the auxiliary oracles (oracle/preprocess_oracle.py, oracle/aux_oracle.py) against the
golden outputs of the reference itself (tests/golden/make_golden_aux.py)."""
import glob
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import aux_oracle as A
from oracle import preprocess_oracle as P
from oracle.synth import synth_image

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


PRE = sorted(glob.glob(os.path.join(GOLD, "preprocess_*.npz")))


def test_preprocess_fixtures_present():
    assert len(PRE) >= 8


@pytest.mark.parametrize("path", PRE, ids=[os.path.basename(p)[11:-4] for p in PRE])
def test_preprocess_oracle_is_bit_exact_against_reference(path):
    g = np.load(path)
    H, W = (int(v) for v in g["hw"])
    img = synth_image(H, W, int(g["seed"]))
    new = P.get_preprocess_shape(H, W)
    assert tuple(int(v) for v in g["resize"]) == new
    small = P.pil_resize_bilinear(img, *new)
    assert _sha(small) == str(g["sha_u8"])
    x, resize = P.preprocess_frame(img)
    assert resize == new and x.dtype == np.float32
    assert np.array_equal(x.reshape(-1)[g["sample_idx"]], g["sample_val"])
    assert _sha(x) == str(g["sha_f32"])
    assert _sha(torch.from_numpy(x).bfloat16().view(torch.int16).numpy()) == str(g["sha_bf16"])


def test_seg_projection_oracle_against_reference():
    g = np.load(os.path.join(GOLD, "aux_segproj.npz"))
    embs, offsets = A.seg_project(g["hidden"], g["output_ids"], int(g["seg_idx"]), g["w_0_weight"], g["w_0_bias"],
                                  g["w_2_weight"], g["w_2_bias"])
    assert offsets.tolist() == g["offsets"].tolist() == [0, 3, 3, 4]
    got = np.concatenate(embs, 0)
    assert got.shape == g["pred"].shape
    assert np.abs(got - g["pred"]).max() <= 2e-5 * np.abs(g["pred"]).max()
    assert embs[1].shape == (0, 256)          # an image without [SEG] yields an empty prompt list (chat.py:227)
    bi, ri, _ = A.seg_rows(g["output_ids"], int(g["seg_idx"]))
    assert bi.tolist() == [0, 0, 0, 2] and ri.tolist() == [255 + 4, 255 + 16, 255 + 29, 255 + 38]


def test_metrics_oracle_against_reference():
    g = np.load(os.path.join(GOLD, "aux_metrics.npz"))
    for i in range(g["left"].shape[0]):
        pred = A.gated_mask(g["left"][i], g["right"][i], g["tax"][i])
        assert np.array_equal(pred, g["pred"][i])
        iou, iocm = A.iou_iocm(pred, g["gt"][i])
        assert iou == g["iou"][i] and iocm == g["iocm"][i]          # integer counts, one double division: exact
