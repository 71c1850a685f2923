"""Model head specs (from the darknet cfgs the reference ships) and the seeded synthetic
inputs of SURVEY.md §8(d).  Shared by tests/ and bench.py; no oracle import here."""
from dataclasses import dataclass
from typing import List, Tuple

import numpy as np
import torch

SEED0 = 20261104

_V4 = [(12, 16), (19, 36), (40, 28), (36, 75), (76, 55), (72, 146), (142, 110), (192, 243), (459, 401)]
_TINY = [(10, 14), (23, 27), (37, 58), (81, 82), (135, 169), (344, 319)]


@dataclass
class HeadSpec:
    name: str
    image: int
    shapes: List[Tuple[int, int]]
    anchors: List[List[Tuple[float, float]]]  # per layer, ratio (h, w)
    num_classes: int = 80

    @property
    def num_instances(self):
        return sum(len(a) * h * w for (h, w), a in zip(self.shapes, self.anchors))

    def head_shape(self, layer, batch):
        H, W = self.shapes[layer]
        return (batch, len(self.anchors[layer]) * (5 + self.num_classes), H, W)


def _ratio(pairs_wh, masks, image):
    # darknet anchors are (w, h) in pixels; the reference stores ratio (h, w)
    return [[(pairs_wh[m][1] / image, pairs_wh[m][0] / image) for m in mask] for mask in masks]


def yolov4_608() -> HeadSpec:
    """darknet-config/cfg/yolov4.cfg:965-969,1053-1057,1141-1145."""
    return HeadSpec("yolov4-608", 608, [(76, 76), (38, 38), (19, 19)],
                    _ratio(_V4, [(0, 1, 2), (3, 4, 5), (6, 7, 8)], 608))


def yolov4_tiny_416() -> HeadSpec:
    """darknet-config/cfg/yolov4-tiny.cfg:217-221,266-270."""
    return HeadSpec("yolov4-tiny-416", 416, [(13, 13), (26, 26)],
                    _ratio(_TINY, [(3, 4, 5), (1, 2, 3)], 416))


def yolov4_csp_640() -> HeadSpec:
    """darknet-config/cfg/yolov4-csp.cfg:1029-1033,1143-1147,1257-1261 at 640x640."""
    return HeadSpec("yolov4-csp-640", 640, [(80, 80), (40, 40), (20, 20)],
                    _ratio(_V4, [(0, 1, 2), (3, 4, 5), (6, 7, 8)], 640))


def small_spec(num_classes=5) -> HeadSpec:
    """a tiny odd-shaped head for fast tests (non-multiple-of-4 maps, uneven anchors)."""
    return HeadSpec("small", 64, [(8, 12), (5, 7), (3, 3)],
                    [[(0.1, 0.15), (0.2, 0.1), (0.3, 0.33)], [(0.3, 0.5), (0.6, 0.4)],
                     [(0.7, 0.9), (0.5, 0.5), (0.9, 0.6)]], num_classes)


SPECS = {"yolov4-608": yolov4_608, "yolov4-tiny-416": yolov4_tiny_416,
         "yolov4-csp-640": yolov4_csp_640, "small": small_spec}


def make_heads(spec: HeadSpec, batch: int, seed: int, device="cpu", obj_mean=-4.0, cls_mean=-3.0):
    """box channels ~N(0,1), objectness ~N(obj_mean,1.5), classes ~N(cls_mean,1.5).
    Generated with a generator of `device` (bytes are identical for equal device type)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    heads = []
    for l, (H, W) in enumerate(spec.shapes):
        A, E = len(spec.anchors[l]), 5 + spec.num_classes
        t = torch.randn((batch, E, A, H, W), generator=g, dtype=torch.float32, device=device)
        t[:, 4].mul_(1.5).add_(obj_mean)
        t[:, 5:].mul_(1.5).add_(cls_mean)
        heads.append(t.view(batch, E * A, H, W))
    return heads


def make_targets(spec: HeadSpec, batch: int, seed: int, n_min=1, n_max=30, clustered=0.0,
                 adversarial=False):
    """GT lists: n ~U{n_min..n_max}; cy,cx ~U(.05,.95); h,w log-U(.02,.6) clipped to the image;
    class ~U{0..C-1}.  `clustered`: fraction of boxes re-centred within 2 cells (finest map) of
    an earlier box to force slot conflicts; `adversarial`: adds centres at exact k and k+0.5 grid
    positions of every layer."""
    rng = np.random.default_rng(seed)
    H0, W0 = spec.shapes[0]
    out = []
    for _ in range(batch):
        n = int(rng.integers(n_min, n_max + 1))
        cy = rng.uniform(0.05, 0.95, n)
        cx = rng.uniform(0.05, 0.95, n)
        h = np.exp(rng.uniform(np.log(0.02), np.log(0.6), n))
        w = np.exp(rng.uniform(np.log(0.02), np.log(0.6), n))
        if clustered > 0 and n > 1:
            for i in range(1, n):
                if rng.random() < clustered:
                    j = int(rng.integers(0, i))
                    cy[i] = np.clip(cy[j] + rng.uniform(-2, 2) / H0, 0.01, 0.99)
                    cx[i] = np.clip(cx[j] + rng.uniform(-2, 2) / W0, 0.01, 0.99)
        h = np.minimum(h, 2 * np.minimum(cy, 1 - cy))
        w = np.minimum(w, 2 * np.minimum(cx, 1 - cx))
        cls = rng.integers(0, spec.num_classes, n)
        boxes = [(float(a), float(b), float(c), float(d), int(e)) for a, b, c, d, e in zip(cy, cx, h, w, cls)]
        if adversarial:
            for (H, W) in spec.shapes:
                k = int(rng.integers(1, H - 1))
                m = int(rng.integers(1, W - 1))
                boxes.append((k / H, (m + 0.5) / W, 0.1, 0.1, 0))
                boxes.append(((k + 0.5) / H, m / W, 0.1, 0.1, 1 % spec.num_classes))
                boxes.append(((k + 0.5) / H, (m + 0.5) / W, 0.12, 0.1, 2 % spec.num_classes))
                boxes.append(((k + 0.5) / H, (m + 0.5) / W, 0.1, 0.12, 3 % spec.num_classes))  # exact tie
            boxes.append((0.5, 0.5, 0.0, 0.2, 0))  # zero-sized: ignored (pred_target_matching.rs:62-73)
        out.append(boxes)
    return out


def make_nms_boxes(K: int, seed: int, num_groups: int = 1, dup_frac: float = 0.05, thr: float = 0.1):
    """planted NMS candidates (SURVEY §8d cfg 4): centres U(0,1), sizes log-U(.01,.3), scores
    U(thr,1) with `dup_frac` exact duplicates, groups U{0..num_groups-1}.
    Returns ltrb [K,4] f32, scores [K] f32, groups [K] i64 (numpy)."""
    rng = np.random.default_rng(seed)
    cy, cx = rng.uniform(0, 1, K), rng.uniform(0, 1, K)
    h = np.exp(rng.uniform(np.log(0.01), np.log(0.3), K))
    w = np.exp(rng.uniform(np.log(0.01), np.log(0.3), K))
    ltrb = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1).astype(np.float32)
    scores = rng.uniform(thr, 1.0, K).astype(np.float32)
    nd = int(K * dup_frac)
    if nd and K > 1:
        src = rng.integers(0, K, nd)
        dst = rng.integers(0, K, nd)
        scores[dst] = scores[src]
    groups = rng.integers(0, num_groups, K).astype(np.int64)
    return ltrb, scores, groups
