"""Generates tests/golden/hotpath_small.npz from the oracle (run in the build container:
`python tests/golden/make_golden.py`).

The reference (Rust/tch) cannot run here and ships no vectors for this path, so these are NOT
reference outputs: they freeze the oracle's answers (torch 2.11 CPU ATen kernels) for a small
head so that drift of the oracle or of the CUDA path is caught.  Inputs are stored too — the
file does not depend on torch's RNG stream.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import detect as odet, loss as oloss, nms as onms  # noqa: E402
from yolo_dl_b200 import synth  # noqa: E402


def main():
    spec = synth.small_spec()
    B = 2
    heads = synth.make_heads(spec, B, synth.SEED0, obj_mean=-0.5, cls_mean=-1.0)
    targets = synth.make_targets(spec, B, synth.SEED0, n_max=6, clustered=0.3, adversarial=True)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    out = {}
    for l, h in enumerate(heads):
        out[f"head{l}"] = h.numpy()
    gt = np.array([(b,) + tuple(r) for b, t in enumerate(targets) for r in t], dtype=np.float64)
    out["gt"] = gt  # rows: batch, cy, cx, h, w, class
    merged = odet.forward_merged(heads, infos, spec.num_classes)
    for k in ("cy", "cx", "h", "w"):
        out[f"dec_{k}"] = merged[k].numpy()
    for metric in ("iou", "giou", "diou", "ciou"):
        r = oloss.yolo_loss(heads, infos, spec.num_classes, targets, oloss.LossConfig(box_metric=metric))
        out[f"{metric}_losses"] = np.array([float(r["losses"][k]) for k in
                                            ("total", "iou", "classification", "objectness")], dtype=np.float32)
        out[f"{metric}_iou_score"] = r["iou_score"].numpy()
        for l, g in enumerate(r["grads"]):
            out[f"{metric}_grad{l}"] = g.numpy()
    out["match_batches"], out["match_flats"], out["match_gt"] = r["matches"][0], r["matches"][1], r["matches"][2]
    for by_class in (False, True):
        k = onms.nms_forward(merged, 0.3, 0.5, by_class)
        tag = "cls" if by_class else "img"
        out[f"nms_{tag}_rows"] = np.stack([k["batches"], k["classes"], k["instances"]], 1)
        out[f"nms_{tag}_tlbr"] = k["tlbr"]
        out[f"nms_{tag}_conf"] = k["conf"]
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hotpath_small.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", len(out["match_flats"]), "matches;",
          len(out["nms_img_rows"]), "/", len(out["nms_cls_rows"]), "kept")


if __name__ == "__main__":
    main()
