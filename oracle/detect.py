"""Detect2D / MergeDetect2D / MergedDenseDetection restated with the reference's ATen op order.

Test infrastructure only (see oracle/__init__.py).
"""
from dataclasses import dataclass
from typing import List, Sequence, Tuple

import torch


@dataclass
class LayerInfo:
    """DetectionInfo (tch-goodies/src/detection/merged_dense_detection.rs:497-507)."""

    H: int
    W: int
    anchors: List[Tuple[float, float]]  # ratio (h, w), f64 as the reference stores them
    flat_start: int = 0

    @property
    def A(self):
        return len(self.anchors)

    @property
    def size(self):
        return self.A * self.H * self.W


def make_infos(shapes: Sequence[Tuple[int, int]], anchors) -> List[LayerInfo]:
    """flat ranges are consecutive (merged_dense_detection.rs:68-72)."""
    infos, start = [], 0
    for (H, W), a in zip(shapes, anchors):
        infos.append(LayerInfo(H, W, [tuple(map(float, x)) for x in a], start))
        start += infos[-1].size
    return infos


def detect2d_forward(head: torch.Tensor, num_classes: int, anchors, divisor_as_tensor: bool = False):
    """Detect2D::forward (tch-modules/src/detect_2d.rs:42-152), cache (:196-221).

    head: [B, A*(5+C), H, W] f32.  Returns dict of cy,cx,h,w [B,1,A,H,W] and the
    obj_logit / class_logit VIEWS of the input.

    `tensor / f64 scalar` is what the reference writes (:69-70,:128-131).  ATen evaluates it as an
    IEEE division on the CPU but as a multiplication by the f32 reciprocal on CUDA
    (BinaryDivTrueKernel.cu, "may lose one bit of precision"), so the reference's own CPU and CUDA
    results differ in the last bit of cy / cx.  `divisor_as_tensor` divides by a 0-dim tensor ON THE
    DEVICE instead, which takes ATen's true-division kernel on either device: the CUDA executor
    then follows the CPU's (and this library's) arithmetic bit for bit.
    """
    B, ch, H, W = head.shape
    A = len(anchors)
    E = num_classes + 5
    assert ch == A * E
    # cache (:196-221): arange/size in f32; anchors cast R64 -> f32
    y_offsets = torch.arange(H, dtype=torch.float32, device=head.device) / float(H)
    x_offsets = torch.arange(W, dtype=torch.float32, device=head.device) / float(W)
    anchor_h = torch.tensor([float(a[0]) for a in anchors], dtype=torch.float64).to(torch.float32)
    anchor_w = torch.tensor([float(a[1]) for a in anchors], dtype=torch.float64).to(torch.float32)
    anchor_h, anchor_w = anchor_h.to(head.device), anchor_w.to(head.device)

    outputs = head.view(B, E, A, H, W)  # (:66)
    dh, dw = float(H), float(W)
    if divisor_as_tensor:
        dh = torch.tensor(float(H), dtype=torch.float32, device=head.device)
        dw = torch.tensor(float(W), dtype=torch.float32, device=head.device)
        y_offsets = torch.arange(H, dtype=torch.float32, device=head.device) / dh
        x_offsets = torch.arange(W, dtype=torch.float32, device=head.device) / dw
    cy = (outputs[:, 0:1].sigmoid() * 2.0 - 0.5) / dh + y_offsets.view(1, 1, 1, H, 1)
    cx = (outputs[:, 1:2].sigmoid() * 2.0 - 0.5) / dw + x_offsets.view(1, 1, 1, 1, W)
    h = outputs[:, 2:3].sigmoid().mul(2.0).pow(2) * anchor_h.view(1, 1, A, 1, 1)
    w = outputs[:, 3:4].sigmoid().mul(2.0).pow(2) * anchor_w.view(1, 1, A, 1, 1)
    obj_logit = outputs[:, 4:5]
    class_logit = outputs[:, 5:]
    return dict(cy=cy, cx=cx, h=h, w=w, obj_logit=obj_logit, class_logit=class_logit)


def merge(dets: Sequence[dict], num_classes: int):
    """MergedDenseDetection::from_detection_tensors (merged_dense_detection.rs:19-119)."""
    B = dets[0]["cy"].shape[0]
    out = {}
    for k in ("cy", "cx", "h", "w", "obj_logit"):
        out[k] = torch.cat([d[k].reshape(B, 1, -1) for d in dets], 2)
    out["class_logit"] = torch.cat([d["class_logit"].reshape(B, num_classes, -1) for d in dets], 2)
    return out


def forward_merged(heads: Sequence[torch.Tensor], infos: Sequence[LayerInfo], num_classes: int,
                   divisor_as_tensor: bool = False):
    dets = [detect2d_forward(h, num_classes, i.anchors, divisor_as_tensor) for h, i in zip(heads, infos)]
    return merge(dets, num_classes)


def index_by_flats(merged: dict, batches: torch.Tensor, flats: torch.Tensor):
    """MergedDenseDetection::index_by_flats (merged_dense_detection.rs:280-307):
    tensor.index([batches, None, flats]) -> [M, entries]."""
    return {k: v[batches, :, flats] for k, v in merged.items()}


def instance_to_flat(info: LayerInfo, anchor: int, row: int, col: int) -> int:
    """merged_dense_detection.rs:374-376."""
    return info.flat_start + col + info.W * (row + info.H * anchor)
