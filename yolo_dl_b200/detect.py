"""Detect2D / MergeDetect2D / MergedDenseDetection over the C ABI (yh_decode, yh_decode_merged).

Mirrors tch-modules/src/detect_2d.rs:6-152, merge_detect_2d.rs:14-19 and
tch-goodies/src/detection/merged_dense_detection.rs:19-119.
"""
from typing import List, Sequence, Tuple

import torch

from . import _lib
from .tensors import DenseDetectionTensor, DenseDetectionTensorList, DetectionInfo


def _stream(t: torch.Tensor):
    return torch.cuda.current_stream(t.device).cuda_stream


def _require_cuda_f32(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise _lib.YoloHotError(_lib.YH_ERR_CUDA, f"{what} is not a CUDA tensor; yolohot has no CPU fallback")
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise ValueError(f"{what} must be a contiguous float32 tensor")


class Detect2DInit:
    """detect_2d.rs:6-31."""

    def __init__(self, num_classes: int, anchors: Sequence[Tuple[float, float]]):
        self.num_classes = int(num_classes)
        self.anchors = [(float(h), float(w)) for h, w in anchors]  # Ratio<HW<R64>>

    def build(self) -> "Detect2D":
        return Detect2D(self.num_classes, self.anchors)


class Detect2D:
    def __init__(self, num_classes, anchors):
        self.num_classes = num_classes
        self.anchors = anchors
        self._lib = _lib.load()

    def info(self, H, W, flat_start=0) -> DetectionInfo:
        return DetectionInfo(H, W, list(self.anchors), flat_start)

    def forward(self, tensor: torch.Tensor) -> DenseDetectionTensor:
        """detect_2d.rs:42-152: `[B, A*(5+C), H, W]` -> DenseDetectionTensor; obj/class logits
        stay views of the input (:136-139)."""
        _require_cuda_f32(tensor, "Detect2D input")
        B, ch, H, W = tensor.shape
        A, E = len(self.anchors), self.num_classes + 5
        if ch != A * E:
            raise ValueError(f"channels {ch} != anchors {A} x entries {E}")
        out = torch.empty((4, B, 1, A, H, W), dtype=torch.float32, device=tensor.device)
        layer = _lib.make_layers([self.info(H, W)])
        with torch.cuda.device(tensor.device):
            _lib.check(self._lib.yh_decode(tensor.data_ptr(), B, self.num_classes, layer,
                                           out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                           out[3].data_ptr(), A * H * W, _stream(tensor)))
        v = tensor.view(B, E, A, H, W)
        return DenseDetectionTensor(out[0], out[1], out[2], out[3], v[:, 4:5], v[:, 5:],
                                    list(self.anchors), tensor)


class MergeDetect2D:
    """merge_detect_2d.rs:14-19 (no math)."""

    def forward(self, tensors: Sequence[DenseDetectionTensor]) -> DenseDetectionTensorList:
        return DenseDetectionTensorList.from_detection_tensors(tensors)


class MergedDenseDetection:
    """merged_dense_detection.rs:465-507: cy,cx,h,w `[B,1,N]`, obj_logit `[B,1,N]`, class_logit
    `[B,C,N]`, info.

    B200 layout: the box planes are ONE `[4,B,1,N]` buffer written by the decode kernel; the
    logits are never copied — `heads` keeps the L head tensors and the loss / NMS kernels read
    them in place (flat index -> (layer, anchor, cell) on the fly).  `obj_logit` / `class_logit`
    materialise the reference's `cat` only if somebody asks for them.
    """

    def __init__(self, boxes: torch.Tensor, heads: List[torch.Tensor], info: List[DetectionInfo],
                 num_classes: int):
        self.boxes = boxes  # [4,B,1,N]
        self.heads = heads
        self.info = info
        self._num_classes = num_classes

    cy = property(lambda s: s.boxes[0])
    cx = property(lambda s: s.boxes[1])
    h = property(lambda s: s.boxes[2])
    w = property(lambda s: s.boxes[3])

    def device(self):
        return self.boxes.device

    def batch_size(self):
        return self.boxes.shape[1]

    def num_instances(self):
        return self.boxes.shape[3]

    def num_classes(self):
        return self._num_classes

    def _entries(self, lo, hi):
        B, E = self.batch_size(), self._num_classes + 5
        parts = [hd.view(B, E, i.A, i.H, i.W)[:, lo:hi].reshape(B, hi - lo, -1)
                 for hd, i in zip(self.heads, self.info)]
        return torch.cat(parts, 2)

    @property
    def obj_logit(self):
        return self._entries(4, 5)

    @property
    def class_logit(self):
        return self._entries(5, self._num_classes + 5)

    @staticmethod
    def make_info(shapes, anchors) -> List[DetectionInfo]:
        infos, start = [], 0
        for (H, W), a in zip(shapes, anchors):
            infos.append(DetectionInfo(H, W, [(float(h), float(w)) for h, w in a], start))
            start += infos[-1].size
        return infos

    @staticmethod
    def from_detection_tensors(tensors) -> "MergedDenseDetection":
        """merged_dense_detection.rs:19-119 — drop-in form: the per-layer decoded planes already
        exist, so they are concatenated (one copy of 4*N*B floats, not of the whole head)."""
        if isinstance(tensors, DenseDetectionTensorList):
            tensors = tensors.tensors
        tensors = list(tensors)
        if len({t.batch_size() for t in tensors}) != 1 or len({t.num_classes() for t in tensors}) != 1:
            raise ValueError("batch size / class count mismatch")
        B = tensors[0].batch_size()
        info = MergedDenseDetection.make_info([(t.height(), t.width()) for t in tensors],
                                              [t.anchors for t in tensors])
        boxes = torch.stack([torch.cat([getattr(t, k).reshape(B, 1, -1) for t in tensors], 2)
                             for k in ("cy", "cx", "h", "w")])
        return MergedDenseDetection(boxes, [t.head for t in tensors], info, tensors[0].num_classes())

    @staticmethod
    def from_heads(heads: Sequence[torch.Tensor], detects: Sequence[Detect2D]) -> "MergedDenseDetection":
        """Fused Detect2D x L + merge: one launch writes all layers straight into
        the merged `[B,1,N]` planes."""
        heads = list(heads)
        for i, hd in enumerate(heads):
            _require_cuda_f32(hd, f"head {i}")
        C = detects[0].num_classes
        B = heads[0].shape[0]
        info = MergedDenseDetection.make_info([tuple(hd.shape[2:]) for hd in heads],
                                              [d.anchors for d in detects])
        for hd, i in zip(heads, info):
            if hd.shape[0] != B or hd.shape[1] != i.A * (C + 5):
                raise ValueError("head shape mismatch")
        N = sum(i.size for i in info)
        boxes = torch.empty((4, B, 1, N), dtype=torch.float32, device=heads[0].device)
        lib = _lib.load()
        with torch.cuda.device(heads[0].device):
            _lib.check(lib.yh_decode_merged(_lib.ptr_array(heads), _lib.make_layers(info), len(info),
                                            B, C, boxes[0].data_ptr(), boxes[1].data_ptr(),
                                            boxes[2].data_ptr(), boxes[3].data_ptr(), N,
                                            _stream(heads[0])))
        return MergedDenseDetection(boxes, heads, info, C)
