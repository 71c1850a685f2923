"""Host-side mirrors of the reference's hot-path data types.

The reference's host language is Rust (no cargo/rustc in this image), so the operator
interface is mirrored in Python over the C ABI with the reference's names and field meaning;
the Rust binding a maintainer would add is in INTEGRATION.md.  torch is used for device memory
and streams only.
"""
import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import torch


@dataclass
class RatioLabel:
    """`RatioLabel = Ratio<Label<CyCxHW<R64>, usize>>` (yolo-dl/src/label.rs:6,
    label/src/lib.rs:6-12): box centre/size in ratio units of the image (f64) + class."""

    cy: float
    cx: float
    h: float
    w: float
    class_: int

    def astuple(self):
        return (float(self.cy), float(self.cx), float(self.h), float(self.w), int(self.class_))


def label_row(x) -> Tuple[float, float, float, float, int]:
    if isinstance(x, RatioLabel):
        return x.astuple()
    cy, cx, h, w, c = x
    return (float(cy), float(cx), float(h), float(w), int(c))


@dataclass
class DetectionInfo:
    """tch-goodies/src/detection/merged_dense_detection.rs:497-507."""

    H: int
    W: int
    anchors: List[Tuple[float, float]]  # ratio (h, w) f64
    flat_start: int = 0

    @property
    def A(self):
        return len(self.anchors)

    @property
    def size(self):
        return self.A * self.H * self.W

    @property
    def flat_index_range(self):
        return range(self.flat_start, self.flat_start + self.size)


@dataclass
class DenseDetectionTensor:
    """tch-goodies/src/detection/dense_detection_tensor.rs:551-575: cy,cx,h,w `[B,1,A,H,W]`,
    obj_logit `[B,1,A,H,W]` and class_logit `[B,C,A,H,W]` (views of the head), anchors."""

    cy: torch.Tensor
    cx: torch.Tensor
    h: torch.Tensor
    w: torch.Tensor
    obj_logit: torch.Tensor
    class_logit: torch.Tensor
    anchors: List[Tuple[float, float]]
    head: torch.Tensor = None  # the [B, A*(5+C), H, W] tensor the views come from

    def batch_size(self):
        return self.cy.shape[0]

    def num_classes(self):
        return self.class_logit.shape[1]

    def num_anchors(self):
        return self.cy.shape[2]

    def height(self):
        return self.cy.shape[3]

    def width(self):
        return self.cy.shape[4]

    # -- the container operations of dense_detection_tensor.rs (host glue: plain tensor ops on
    #    whatever device the fields live on; pinned by the reference's own `cat_test`, :582-728,
    #    in tests/test_tensors_known_answers.py) -------------------------------------------
    def check(self) -> "DenseDetectionTensor":
        """TryFrom<DenseDetectionTensorUnchecked> (:551-575)"""
        if self.class_logit.dim() != 5:
            raise ValueError("class_logit must have 5 dimensions")
        B, _, A, H, W = self.class_logit.shape
        for name in ("cy", "cx", "h", "w", "obj_logit"):
            if tuple(getattr(self, name).shape) != (B, 1, A, H, W):
                raise ValueError(f"{name} must have shape {(B, 1, A, H, W)}")
        if len(self.anchors) != A:
            raise ValueError("the number of anchors does not match the tensors")
        return self

    def obj_prob(self):
        return self.obj_logit.sigmoid()  # :42-44

    def class_prob(self):
        return self.class_logit.sigmoid()  # :46-48

    def confidence(self):
        return self.obj_prob() * self.class_prob()  # :51-53

    def unbiased_cy(self):
        H = self.height()  # :55-62
        off = (torch.arange(H, dtype=torch.float32, device=self.cy.device) / float(H)).view(1, 1, 1, H, 1)
        return self.cy - off

    def unbiased_cx(self):
        W = self.width()  # :64-71
        off = (torch.arange(W, dtype=torch.float32, device=self.cx.device) / float(W)).view(1, 1, 1, 1, W)
        return self.cx - off

    def to_tensor(self):
        return torch.cat([self.cy, self.cx, self.h, self.w, self.obj_logit, self.class_logit], 1)  # :207-218

    def _fields(self):
        return [self.cy, self.cx, self.h, self.w, self.obj_logit, self.class_logit]

    def index_select_batch(self, index: torch.Tensor) -> "DenseDetectionTensor":
        """:73-108"""
        if index.dim() != 1 or index.dtype != torch.int64:
            raise ValueError("index must be a 1-D int64 tensor")
        if not bool((index < self.batch_size()).all()):
            raise ValueError(f"batch index exceeds batch size {self.batch_size()}")
        return DenseDetectionTensor(*[t.index_select(0, index) for t in self._fields()], list(self.anchors))

    def slice_ratio(self, y_range: Tuple[float, float], x_range: Tuple[float, float]) -> "DenseDetectionTensor":
        """:110-131 (f64::round = half away from zero)"""
        ys, ye = y_range
        xs, xe = x_range
        if not (0.0 <= ys <= 1.0 and 0.0 <= ye <= 1.0 and 0.0 <= xs <= 1.0 and 0.0 <= xe <= 1.0
                and ye - ys > 0.0 and xe - xs > 0.0):
            raise ValueError("invalid ratio range")
        rnd = lambda v: int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)
        H, W = self.height(), self.width()
        new_h, new_w = rnd(H * (ye - ys)), rnd(W * (xe - xs))
        y0, x0 = rnd(ys * H), rnd(xs * W)
        return self.slice((y0, y0 + new_h), (x0, x0 + new_w))

    def slice(self, y_range: Tuple[int, int], x_range: Tuple[int, int]) -> "DenseDetectionTensor":
        """crop to rows/columns [start, end) and re-express boxes and anchors in the units of the
        crop (:133-205)"""
        H, W = self.height(), self.width()
        (y0, y1), (x0, x1) = y_range, x_range
        new_h, new_w = y1 - y0, x1 - x0
        if not (0 <= y0 < H and 1 <= y1 <= H and 0 <= x0 < W and 1 <= x1 <= W and new_h > 0 and new_w > 0):
            raise ValueError("slice range outside the feature map")
        cy, cx, h, w, obj, cls = [t[:, :, :, y0:y1, x0:x1].contiguous() for t in self._fields()]
        cy = (cy * float(H) - float(y0)) / float(new_h)
        cx = (cx * float(W) - float(x0)) / float(new_w)
        h = h * float(H) / float(new_h)
        w = w * float(W) / float(new_w)
        anchors = [(ah * H / new_h, aw * W / new_w) for ah, aw in self.anchors]
        return DenseDetectionTensor(cy, cx, h, w, obj, cls, anchors).check()

    @staticmethod
    def _same_shapes(tensors):
        tensors = list(tensors)
        if not tensors:
            raise ValueError("no tensors")
        for what in ("batch_size", "num_classes", "height", "width"):
            if len({getattr(t, what)() for t in tensors}) != 1:
                raise ValueError(f"{what} mismatch")  # ensure!(.._set.len() == 1)
        if len({tuple(map(tuple, t.anchors)) for t in tensors}) != 1:
            raise ValueError("anchors mismatch")
        return tensors

    @staticmethod
    def cat_batch(tensors) -> "DenseDetectionTensor":
        """:220-304"""
        ts = DenseDetectionTensor._same_shapes(tensors)
        cols = zip(*[t._fields() for t in ts])
        return DenseDetectionTensor(*[torch.cat(list(c), 0) for c in cols], list(ts[0].anchors))

    @staticmethod
    def cat_height(tensors) -> "DenseDetectionTensor":
        """stack feature maps top to bottom: cy -> (cy + i)/n, h -> h/n, anchors h/n (:306-404)"""
        ts = DenseDetectionTensor._same_shapes(tensors)
        n = float(len(ts))
        cy = torch.cat([(t.cy + float(i)) / n for i, t in enumerate(ts)], 3)
        h = torch.cat([t.h / n for t in ts], 3)
        rest = {k: torch.cat([getattr(t, k) for t in ts], 3) for k in ("cx", "w", "obj_logit", "class_logit")}
        anchors = [(ah / n, aw) for ah, aw in ts[0].anchors]
        return DenseDetectionTensor(cy, rest["cx"], h, rest["w"], rest["obj_logit"], rest["class_logit"], anchors)

    @staticmethod
    def cat_width(tensors) -> "DenseDetectionTensor":
        """stack feature maps left to right: cx -> (cx + i)/n, w -> w/n, anchors w/n (:406-504)"""
        ts = DenseDetectionTensor._same_shapes(tensors)
        n = float(len(ts))
        cx = torch.cat([(t.cx + float(i)) / n for i, t in enumerate(ts)], 4)
        w = torch.cat([t.w / n for t in ts], 4)
        rest = {k: torch.cat([getattr(t, k) for t in ts], 4) for k in ("cy", "h", "obj_logit", "class_logit")}
        anchors = [(ah, aw / n) for ah, aw in ts[0].anchors]
        return DenseDetectionTensor(rest["cy"], cx, rest["h"], w, rest["obj_logit"], rest["class_logit"], anchors)


@dataclass
class DenseDetectionTensorList:
    """tch-goodies/src/detection/dense_detection_tensor_list.rs:10-28."""

    tensors: List[DenseDetectionTensor]

    @staticmethod
    def from_detection_tensors(tensors: Sequence[DenseDetectionTensor]):
        tensors = list(tensors)
        if len({t.batch_size() for t in tensors}) != 1:
            raise ValueError("batch size mismatch")  # ensure! in the reference
        return DenseDetectionTensorList(tensors)


@dataclass
class FlatIndexTensor:
    """tch-goodies/src/detection/flat_index_tensor.rs:11-42."""

    batches: torch.Tensor  # [M] i64
    flats: torch.Tensor    # [M] i64

    def num_samples(self):
        return self.batches.shape[0]


@dataclass
class CyCxHWTensor:
    cy: torch.Tensor  # [M,1]
    cx: torch.Tensor
    h: torch.Tensor
    w: torch.Tensor


@dataclass
class TLBRTensor:
    t: torch.Tensor  # [K,1]
    l: torch.Tensor
    b: torch.Tensor
    r: torch.Tensor


@dataclass
class ObjectDetectionTensor:
    """tch-goodies/src/detection/object_detection.rs:6-17."""

    cycxhw: CyCxHWTensor
    obj_logit: torch.Tensor    # [M,1]
    class_logit: torch.Tensor  # [M,C]


@dataclass
class LabelTensor:
    """tch-goodies/src/compound_tensor/label.rs:5-11."""

    cycxhw: CyCxHWTensor
    class_: torch.Tensor  # [M,1] i64


@dataclass
class MatchingOutput:
    """yolo-dl/src/loss/pred_target_matching.rs:280-312. Rows sorted by (batch, flat)."""

    pred_indexes: FlatIndexTensor
    pred: ObjectDetectionTensor
    target: LabelTensor
    gt_rows: torch.Tensor = None  # [M] i32: row of the packed GT list that won the slot
    pred_rows: torch.Tensor = None  # [M, 5+C] contiguous rows `pred` is a view of (yh_benchmark input)

    def num_samples(self):
        return self.pred_indexes.num_samples()


@dataclass
class YoloLossOutput:
    """yolo-dl/src/loss/loss_.rs:519-524."""

    total_loss: torch.Tensor
    iou_loss: torch.Tensor
    classification_loss: torch.Tensor
    objectness_loss: torch.Tensor


@dataclass
class YoloLossAuxiliary:
    """yolo-dl/src/loss/loss_.rs:472-475."""

    matchings: MatchingOutput
    iou_score: Optional[torch.Tensor]  # [M,1]


@dataclass
class NmsOutput:
    """yolo-dl/src/loss/nms.rs:50-56 (also YoloInferenceOutput, inference.rs:97-141)."""

    batches: torch.Tensor     # [K] i64
    classes: torch.Tensor     # [K] i64
    instances: torch.Tensor   # [K] i64
    bbox: TLBRTensor          # [K,1] each
    confidence: torch.Tensor  # [K,1]

    def num_samples(self):
        return self.batches.shape[0]

    def index_select(self, indexes):
        b = self.bbox
        return NmsOutput(self.batches[indexes], self.classes[indexes], self.instances[indexes],
                         TLBRTensor(b.t[indexes], b.l[indexes], b.b[indexes], b.r[indexes]),
                         self.confidence[indexes])

    def batch_select(self, batch_index: int):
        """inference.rs:137-140."""
        return self.index_select((self.batches == batch_index).nonzero().view(-1))


YoloInferenceOutput = NmsOutput
