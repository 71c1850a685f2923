"""CyCxHWMatcher / YoloLoss over the C ABI (yh_assign, yh_loss_fwd_bwd).

Mirrors yolo-dl/src/loss/pred_target_matching.rs:11-48 and yolo-dl/src/loss/loss_.rs:21-36,
169-189, 227-277 (names, defaults, validation messages).
"""
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from .detect import MergedDenseDetection, _stream
from .tensors import (CyCxHWTensor, FlatIndexTensor, LabelTensor, MatchingOutput,
                      ObjectDetectionTensor, YoloLossAuxiliary, YoloLossOutput, label_row)
from .workspace import Workspace


def pack_labels(target: Sequence[Sequence]):
    """&[Vec<RatioLabel>] -> ([G,5] f64 rows cy,cx,h,w,class ; [B+1] i32 offsets), pinned."""
    offs = np.zeros(len(target) + 1, dtype=np.int32)
    rows = []
    for i, boxes in enumerate(target):
        offs[i + 1] = offs[i] + len(boxes)
        rows.extend(label_row(x) for x in boxes)
    gt = np.asarray(rows, dtype=np.float64).reshape(-1, 5)
    gt_t = torch.from_numpy(np.ascontiguousarray(gt))
    off_t = torch.from_numpy(offs)
    if torch.cuda.is_available():
        gt_t, off_t = gt_t.pin_memory(), off_t.pin_memory()
    return gt_t, off_t


@dataclass
class CyCxHWMatcherInit:
    """pred_target_matching.rs:11-24."""

    match_grid_method: str = "Rect4"  # MatchGrid::{Rect2, Rect4} (misc.rs:6-12)
    anchor_scale_thresh: float = 4.0

    def build(self) -> "CyCxHWMatcher":
        if not self.anchor_scale_thresh >= 1.0:
            raise ValueError("anchor_scale_thresh must be greater than or equal to 1")
        if self.match_grid_method not in ("Rect2", "Rect4"):
            raise ValueError(f"unknown match_grid_method {self.match_grid_method}")
        return CyCxHWMatcher(self.match_grid_method, float(self.anchor_scale_thresh))


class _Assigned:
    """raw output buffers of yh_assign (capacity rows; M on the device)."""

    def __init__(self, cap, dev):
        self.cap = cap
        self.batch = torch.empty(cap, dtype=torch.int64, device=dev)
        self.flat = torch.empty(cap, dtype=torch.int64, device=dev)
        self.gt = torch.empty(cap, dtype=torch.int32, device=dev)
        self.target = torch.empty((cap, 4), dtype=torch.float32, device=dev)
        self.cls = torch.empty(cap, dtype=torch.int64, device=dev)
        self.M = torch.zeros(1, dtype=torch.int32, device=dev)


class CyCxHWMatcher:
    def __init__(self, match_grid_method, anchor_scale_thresh):
        self.match_grid_method = match_grid_method
        self.anchor_scale_thresh = anchor_scale_thresh
        self._lib = _lib.load()
        self._ws = Workspace()

    def assign(self, prediction: MergedDenseDetection, target, ws: Optional[Workspace] = None) -> _Assigned:
        """yh_assign; no synchronisation. target = &[Vec<RatioLabel>] or a packed (gt, offsets)."""
        dev = prediction.device()
        B, N = prediction.batch_size(), prediction.num_instances()
        gt_h, off_h = target if isinstance(target, tuple) else pack_labels(target)
        if off_h.shape[0] != B + 1:
            raise ValueError(f"{off_h.shape[0] - 1} label lists for batch size {B}")
        G = gt_h.shape[0]
        gt_d = gt_h.to(dev, non_blocking=True)
        off_d = off_h.to(dev, non_blocking=True)
        cells = 5 if self.match_grid_method == "Rect4" else 3
        cap = max(1, G * sum(i.A for i in prediction.info) * cells)
        out = _Assigned(cap, dev)
        ws = ws or self._ws
        h = ws.ensure(dev, B, N)
        with torch.cuda.device(dev):
            _lib.check(self._lib.yh_assign(
                h, gt_d.data_ptr(), off_d.data_ptr(), B, _lib.make_layers(prediction.info),
                len(prediction.info), 1 if self.match_grid_method == "Rect4" else 0,
                self.anchor_scale_thresh, out.batch.data_ptr(), out.flat.data_ptr(),
                out.gt.data_ptr(), out.target.data_ptr(), out.cls.data_ptr(), out.M.data_ptr(),
                cap, _stream(prediction.boxes)))
        out._keep = (gt_d, off_d)
        return out

    def match_bboxes(self, prediction: MergedDenseDetection, target) -> MatchingOutput:
        """pred_target_matching.rs:51-276."""
        a = self.assign(prediction, target)
        M, _ = self._ws.read_counts(_stream(prediction.boxes), M_dev=a.M)
        return _matching_output(prediction, a, M, None)


def _gather_pred_rows(prediction: MergedDenseDetection, batches, flats):
    """index_by_flats (merged_dense_detection.rs:280-307) with torch indexing (glue, standalone
    matcher only; YoloLoss gets the rows from the loss kernel)."""
    B, E = prediction.batch_size(), prediction.num_classes() + 5
    M = batches.shape[0]
    rows = torch.empty((M, E), dtype=torch.float32, device=prediction.device())
    rows[:, :4] = prediction.boxes[:, batches, 0, flats].t()
    for hd, i in zip(prediction.heads, prediction.info):
        sel = ((flats >= i.flat_start) & (flats < i.flat_start + i.size)).nonzero().view(-1)
        if sel.numel():
            v = hd.view(B, E, i.size)
            rows[sel, 4:] = v[batches[sel], 4:, flats[sel] - i.flat_start]
    return rows


def _matching_output(prediction, a: _Assigned, M: int, pred_rows) -> MatchingOutput:
    batches, flats = a.batch[:M], a.flat[:M]
    if pred_rows is None:
        pred_rows = _gather_pred_rows(prediction, batches, flats)
    pr = pred_rows[:M]
    tg = a.target[:M]
    return MatchingOutput(
        FlatIndexTensor(batches, flats),
        ObjectDetectionTensor(CyCxHWTensor(pr[:, 0:1], pr[:, 1:2], pr[:, 2:3], pr[:, 3:4]),
                              pr[:, 4:5], pr[:, 5:]),
        LabelTensor(CyCxHWTensor(tg[:, 0:1], tg[:, 1:2], tg[:, 2:3], tg[:, 3:4]),
                    a.cls[:M].view(-1, 1)),
        a.gt[:M], pr)


@dataclass
class YoloLossInit:
    """loss_.rs:21-36, defaults :169-189 (BCE objectness / classification kinds)."""

    reduction: str = "mean"
    match_grid_method: str = "Rect4"
    anchor_scale_thresh: float = 4.0
    box_metric: str = "diou"  # BoxMetric::{IoU,GIoU,DIoU,CIoU,Hausdorff} (misc.rs:14-21)
    iou_loss_weight: float = 0.05
    objectness_loss_weight: float = 1.0
    classification_loss_weight: float = 0.58
    objectness_pos_weight: Optional[float] = None
    smooth_classification_coef: float = 0.01
    smooth_objectness_coef: float = 0.0
    focal_loss_gamma: float = 0.0              # loss_.rs:176
    objectness_loss_kind: str = "Bce"          # ObjectnessLossKind::{Bce,Focal,L2} (:191-196)
    classification_loss_kind: str = "Bce"      # ClassificationLossKind::{Bce,Focal,L2,CrossEntropy} (:198-205)
    unit_upstream_grad: bool = True  # both reference call sites do total_loss.backward()

    def build(self) -> "YoloLoss":
        # loss_.rs:64-87
        if not self.focal_loss_gamma >= 0.0:
            raise ValueError("focal_loss_gamma must be non-negative")
        if self.objectness_loss_kind.lower() not in ("bce", "focal", "l2"):
            raise ValueError(f"unknown objectness_loss_kind {self.objectness_loss_kind}")
        if self.classification_loss_kind.lower() not in _lib.LOSS_KINDS:
            raise ValueError(f"unknown classification_loss_kind {self.classification_loss_kind}")
        if not 0.0 <= self.smooth_classification_coef <= 1.0:
            raise ValueError("smooth_classification_coef must be in range [0, 1]")
        if not 0.0 <= self.smooth_objectness_coef <= 1.0:
            raise ValueError("smooth_objectness_coef must be in range [0, 1]")
        for name in ("iou_loss_weight", "objectness_loss_weight", "classification_loss_weight"):
            if not getattr(self, name) >= 0.0:
                raise ValueError(f"{name} must be non-negative")
        if self.box_metric.lower() not in _lib.METRICS:
            raise ValueError(f"unknown box_metric {self.box_metric}")
        if self.reduction not in _lib.REDUCTIONS:
            raise ValueError(f"reduction {self.reduction} is not supported")
        matcher = CyCxHWMatcherInit(self.match_grid_method, self.anchor_scale_thresh).build()
        return YoloLoss(self, matcher)


class _HeadGrad(torch.autograd.Function):
    """Delivers the head gradients the loss kernel already wrote when the caller runs
    `total_loss.backward()` (train/src/train/single_gpu.rs:161-168)."""

    @staticmethod
    def forward(ctx, total, unit, *heads):
        ctx.unit = unit
        return total.clone()

    @staticmethod
    def backward(ctx, g):
        grads = ctx.saved_grads
        if ctx.unit:
            torch._assert_async((g == 1.0).all())
            return (None, None, *grads)
        return (None, None, *[x.mul_(g) for x in grads])


class YoloLoss:
    def __init__(self, init: YoloLossInit, matcher: CyCxHWMatcher):
        self.init = init
        self.bbox_matcher = matcher
        self._lib = _lib.load()
        self._ws = Workspace()
        c = _lib.YhLossCfg()
        c.box_metric = _lib.METRICS[init.box_metric.lower()]
        c.reduction = _lib.REDUCTIONS[init.reduction]
        c.iou_loss_weight = init.iou_loss_weight
        c.classification_loss_weight = init.classification_loss_weight
        c.objectness_loss_weight = init.objectness_loss_weight
        c.smooth_classification_coef = init.smooth_classification_coef
        c.smooth_objectness_coef = init.smooth_objectness_coef
        c.has_objectness_pos_weight = 0 if init.objectness_pos_weight is None else 1
        c.objectness_pos_weight = init.objectness_pos_weight or 0.0
        c.objectness_loss_kind = _lib.LOSS_KINDS[init.objectness_loss_kind.lower()]
        c.classification_loss_kind = _lib.LOSS_KINDS[init.classification_loss_kind.lower()]
        c.focal_loss_gamma = init.focal_loss_gamma
        self._cfg = c

    def forward_raw(self, prediction: MergedDenseDetection, target, want_grad: bool):
        """assign + loss (+ head gradients) without any synchronisation.
        Returns (assigned, losses4 [4], iou_score [cap], pred_rows [cap,E], grads | None)."""
        dev = prediction.device()
        B, C = prediction.batch_size(), prediction.num_classes()
        self._ws.ensure(dev, B, prediction.num_instances())
        a = self.bbox_matcher.assign(prediction, target, self._ws)
        losses4 = torch.empty(4, dtype=torch.float32, device=dev)
        iou_score = torch.empty(a.cap, dtype=torch.float32, device=dev)
        pred_rows = torch.empty((a.cap, C + 5), dtype=torch.float32, device=dev)
        grads = [torch.empty_like(h) for h in prediction.heads] if want_grad else None
        if want_grad:
            self._ws.reserve_rows(a.cap, C)
        with torch.cuda.device(dev):
            _lib.check(self._lib.yh_loss_fwd_bwd(
                self._ws.handle, _lib.ptr_array(prediction.heads),
                _lib.make_layers(prediction.info), len(prediction.info), B, C,
                a.batch.data_ptr(), a.flat.data_ptr(), a.target.data_ptr(), a.cls.data_ptr(),
                a.M.data_ptr(), a.cap, self._cfg, losses4.data_ptr(), iou_score.data_ptr(),
                pred_rows.data_ptr(), _lib.ptr_array(grads) if want_grad else None,
                _stream(prediction.boxes)))
        return a, losses4, iou_score, pred_rows, grads

    def forward(self, prediction: MergedDenseDetection, target):
        """loss_.rs:227-277 -> (YoloLossOutput, YoloLossAuxiliary)."""
        want_grad = torch.is_grad_enabled() and any(h.requires_grad for h in prediction.heads)
        a, losses4, iou_score, pred_rows, grads = self.forward_raw(prediction, target, want_grad)
        M, _ = self._ws.read_counts(_stream(prediction.boxes), M_dev=a.M)
        total = losses4[0]
        if want_grad:
            total = _HeadGrad.apply(total, self.init.unit_upstream_grad, *prediction.heads)
            total.grad_fn.saved_grads = [g if h.requires_grad else None
                                         for g, h in zip(grads, prediction.heads)]
        out = YoloLossOutput(total, losses4[1], losses4[2], losses4[3])
        # BoxMetric::Hausdorff yields no score (loss_.rs:285-288)
        haus = self.init.box_metric.lower() == "hausdorff"
        aux = YoloLossAuxiliary(_matching_output(prediction, a, M, pred_rows),
                                None if haus else iou_score[:M].view(-1, 1))
        self.last_grads = grads
        return out, aux
