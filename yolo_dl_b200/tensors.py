"""Host-side mirrors of the reference's hot-path data types.

The reference's host language is Rust (no cargo/rustc in this image), so the operator
interface is mirrored in Python over the C ABI with the reference's names and field meaning;
the Rust binding a maintainer would add is in INTEGRATION.md.  torch is used for device memory
and streams only.
"""
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
