"""GPU parity against the committed golden vectors (tests/golden/hotpath_small.npz, frozen from
the oracle by tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh

from test_oracle import golden_inputs
from util import build, close, grad_close

pytestmark = pytest.mark.gpu


def test_cuda_path_reproduces_golden(dev):
    g, spec, heads_cpu, targets = golden_inputs()
    for metric in ("iou", "giou", "diou", "ciou"):
        heads, _, merged, _ = build(spec, heads_cpu, dev, requires_grad=True)
        for k in ("cy", "cx", "h", "w"):
            close(getattr(merged, k), g[f"dec_{k}"], what=f"golden {k}")
        out, aux = yh.YoloLossInit(box_metric=metric).build().forward(merged, targets)
        out.total_loss.backward()
        got = torch.stack([out.total_loss.detach(), out.iou_loss, out.classification_loss, out.objectness_loss])
        close(got, g[f"{metric}_losses"], rtol=1e-5, what=f"golden {metric} losses")
        assert np.array_equal(aux.matchings.pred_indexes.batches.cpu().numpy(), g["match_batches"])
        assert np.array_equal(aux.matchings.pred_indexes.flats.cpu().numpy(), g["match_flats"])
        for l, h in enumerate(heads):
            grad_close(h.grad, g[f"{metric}_grad{l}"], what=f"golden {metric} grad{l}")
    for by_class, tag in ((False, "img"), (True, "cls")):
        out = yh.NonMaxSuppressionInit(0.5, 0.3, by_class).build().forward(merged)
        rows = torch.stack([out.batches, out.classes, out.instances], 1).cpu().numpy()
        assert set(map(tuple, rows.tolist())) == set(map(tuple, g[f"nms_{tag}_rows"].tolist()))
