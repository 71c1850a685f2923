"""The oracle's box metrics (oracle/box.py = tch-goodies cycxhw.rs:67-120) against the reference
repository's vendored C implementation (darknet box.c:135-256), through committed outputs of that
code (tests/golden/darknet_box.npz, made by tests/golden/make_darknet_box_golden.py) and, where
oracle/_ref is built, through the library itself.

IoU, GIoU and DIoU are the same formulas in both (epsilons aside); CIoU is NOT: yolo-dl ADDS
alpha*v to DIoU (cycxhw.rs:119) where darknet and the paper subtract it — the test pins exactly
that: the two alpha*v terms agree in magnitude and have opposite signs."""
import os

import numpy as np
import pytest
import torch

from oracle import box as obox, build as obuild

HERE = os.path.dirname(os.path.abspath(__file__))


def _golden():
    return np.load(os.path.join(HERE, "golden", "darknet_box.npz"))


def _oracle(a, b):
    # darknet (x, y, w, h) -> (cy, cx, h, w) columns as [M,1] f32 tensors
    col = lambda m, i: torch.from_numpy(m[:, i:i + 1].copy())
    p = (col(a, 1), col(a, 0), col(a, 3), col(a, 2))
    q = (col(b, 1), col(b, 0), col(b, 3), col(b, 2))
    return {k: f(p, q).view(-1).numpy() for k, f in obox.METRICS.items()}


def test_oracle_box_metrics_match_the_reference_trees_c_code():
    g = _golden()
    o = _oracle(g["a_xywh"], g["b_xywh"])
    assert len(g["iou"]) == 4096 and int((g["iou"] > 0).sum()) > 2000 and int((g["iou"] == 0).sum()) > 500
    # f32 evaluation of the same formulas in a different association: a few ulp of values <= 2
    for k in ("iou", "giou", "diou"):
        err = np.abs(o[k].astype(np.float64) - g[k].astype(np.float64))
        assert err.max() <= 2e-6, (k, err.max())
    # CIoU: term = ciou - diou is +alpha*v in yolo-dl and -alpha*v in darknet
    # (alpha's denominator carries +1e-6 in darknet, nothing in yolo-dl: relative 1e-6/(1-iou+v))
    t_oracle = o["ciou"].astype(np.float64) - o["diou"].astype(np.float64)
    t_dk = g["ciou"].astype(np.float64) - g["diou"].astype(np.float64)
    assert (t_oracle >= -1e-7).all() and (t_dk <= 1e-7).all()
    ok = np.abs(t_oracle + t_dk) <= 1e-3 * np.abs(t_dk) + 2e-6
    assert ok.all(), (int((~ok).sum()), np.abs(t_oracle + t_dk).max())
    assert np.abs(t_dk).max() > 0.01  # the term is exercised


def test_oracle_ref_library_reproduces_its_committed_outputs():
    if obuild.build_ref() is None:
        pytest.skip("oracle/_ref is not built and /root/reference is absent")
    g = _golden()
    got = obuild.darknet_box_metrics(g["a_xywh"], g["b_xywh"])
    for k, v in zip(("iou", "giou", "diou", "ciou"), got):
        assert np.array_equal(v, g[k]), k
