"""NonMaxSuppression / YoloInference / nms_by_scores over the C ABI
(yh_nms, yh_inference_pick, yh_nms_by_scores).

Mirrors yolo-dl/src/loss/nms.rs:11-47,72-170 and yolo-dl/src/loss/inference.rs:6-35,37-93.
"""
from dataclasses import dataclass

import torch

from . import _lib
from .detect import MergedDenseDetection, _stream
from .tensors import NmsOutput, TLBRTensor
from .workspace import Workspace


@dataclass
class NonMaxSuppressionInit:
    """nms.rs:11-26 (defaults iou 0.6, conf 0.1, per image)."""

    iou_threshold: float = 0.6
    confidence_threshold: float = 0.1
    suppress_by_class: bool = False
    # Initial capacity of the per-image candidate list.  The reference takes whatever `nonzero()`
    # returns (nms.rs:97-103); here a batch that outgrows the list is detected at the count read
    # forward() does anyway, the workspace is re-created with the size the filter counted and the
    # call repeated (`grow`), so the capacity is a starting point, not a limit.
    max_candidates_per_image: int = 8192
    max_detections: int = 0                # output capacity; 0 = batch * max_candidates_per_image
    grow: bool = True                      # False: an overflow raises YoloHotError(YH_ERR_CAPACITY)

    def build(self) -> "NonMaxSuppression":
        # nms.rs:35-39
        if not self.iou_threshold >= 0.0:
            raise ValueError("iou_threshold must be non-negative")
        if not self.confidence_threshold >= 0.0:
            raise ValueError("confidence_threshold must be non-negative")
        return NonMaxSuppression(self)


class _RawNms:
    def __init__(self, cap, dev):
        self.cap = cap
        self.kb = torch.empty(cap, dtype=torch.int64, device=dev)
        self.kc = torch.empty(cap, dtype=torch.int64, device=dev)
        self.ki = torch.empty(cap, dtype=torch.int64, device=dev)
        self.tlbr = torch.empty((cap, 4), dtype=torch.float32, device=dev)
        self.conf = torch.empty(cap, dtype=torch.float32, device=dev)
        self.K = torch.zeros(1, dtype=torch.int32, device=dev)

    def output(self, K) -> NmsOutput:
        t = self.tlbr[:K]
        return NmsOutput(self.kb[:K], self.kc[:K], self.ki[:K],
                         TLBRTensor(t[:, 0:1], t[:, 1:2], t[:, 2:3], t[:, 3:4]),
                         self.conf[:K].view(-1, 1))


class NonMaxSuppression:
    def __init__(self, init: NonMaxSuppressionInit):
        self.iou_threshold = float(init.iou_threshold)
        self.confidence_threshold = float(init.confidence_threshold)
        self.suppress_by_class = bool(init.suppress_by_class)
        self.max_candidates_per_image = int(init.max_candidates_per_image)
        self.max_detections = int(init.max_detections)
        self.grow = bool(init.grow)
        self._lib = _lib.load()
        self._ws = Workspace()

    def forward_raw(self, prediction: MergedDenseDetection, pick_best_class: bool = False) -> _RawNms:
        """yh_nms (+ yh_inference_pick); no synchronisation."""
        dev = prediction.device()
        B, C = prediction.batch_size(), prediction.num_classes()
        h = self._ws.ensure(dev, B, prediction.num_instances(), self.max_candidates_per_image)
        cap = self.max_detections or B * self.max_candidates_per_image
        out = _RawNms(cap, dev)
        s = _stream(prediction.boxes)
        with torch.cuda.device(dev):
            _lib.check(self._lib.yh_nms(
                h, _lib.ptr_array(prediction.heads), _lib.make_layers(prediction.info),
                len(prediction.info), B, C, self.confidence_threshold, self.iou_threshold,
                1 if self.suppress_by_class else 0, out.kb.data_ptr(), out.kc.data_ptr(),
                out.ki.data_ptr(), out.tlbr.data_ptr(), out.conf.data_ptr(), out.K.data_ptr(),
                cap, s))
            if pick_best_class:
                _lib.check(self._lib.yh_inference_pick(
                    h, out.kb.data_ptr(), out.kc.data_ptr(), out.ki.data_ptr(),
                    out.tlbr.data_ptr(), out.conf.data_ptr(), out.K.data_ptr(), cap, s))
        return out

    def _run(self, prediction: MergedDenseDetection, pick_best_class: bool):
        """forward_raw + the synchronising count read; on a candidate-list overflow the workspace
        grows to what the filter counted and the call is repeated (count -> allocate -> fill)."""
        while True:
            raw = self.forward_raw(prediction, pick_best_class)
            try:
                _, K = self._ws.read_counts(_stream(prediction.boxes), K_dev=raw.K)
                return raw, K
            except _lib.YoloHotError as e:
                need = int(self._lib.yh_workspace_candidates_needed(self._ws.handle))
                if not (self.grow and e.status == _lib.YH_ERR_CAPACITY and
                        need > self.max_candidates_per_image):
                    raise
                del raw
                cap = 1
                while cap < need:
                    cap <<= 1
                self.max_candidates_per_image = cap

    def forward(self, prediction: MergedDenseDetection) -> NmsOutput:
        """nms.rs:72-170. Kept rows ordered by (batch, confidence desc, class*N+instance)."""
        with torch.no_grad():
            raw, K = self._run(prediction, False)
            return raw.output(K)


@dataclass
class YoloInferenceInit:
    """inference.rs:6-10."""

    nms_iou_thresh: float
    nms_conf_thresh: float
    suppress_by_class: bool
    max_candidates_per_image: int = 8192

    def build(self) -> "YoloInference":
        nms = NonMaxSuppressionInit(self.nms_iou_thresh, self.nms_conf_thresh,
                                    self.suppress_by_class, self.max_candidates_per_image).build()
        return YoloInference(nms)


class YoloInference:
    def __init__(self, nms: NonMaxSuppression):
        self.nms = nms

    def forward(self, prediction: MergedDenseDetection) -> NmsOutput:
        """inference.rs:37-93: NMS, then per (batch, instance) the max-confidence class, on the
        device (the reference round-trips through the host)."""
        with torch.no_grad():
            raw, K = self.nms._run(prediction, True)
            return raw.output(K)


_ws_by_scores = {}


def nms_by_scores(ltrb: torch.Tensor, scores: torch.Tensor, groups: torch.Tensor,
                  iou_threshold: float) -> torch.Tensor:
    """Drop-in for `tch_nms::nms_by_scores` (call site nms.rs:142-148): kept row indices (i64),
    score-descending, ties keep the lower row."""
    if not ltrb.is_cuda:
        raise _lib.YoloHotError(_lib.YH_ERR_CUDA, "nms_by_scores needs CUDA tensors; there is no CPU fallback")
    K = scores.shape[0]
    ltrb = ltrb.contiguous().float()
    scores = scores.contiguous().float().view(-1)
    groups = groups.contiguous().long().view(-1)
    if ltrb.shape != (K, 4) or groups.shape[0] != K:
        raise ValueError("size mismatch")
    dev = ltrb.device
    ws = _ws_by_scores.setdefault(dev.index, Workspace())
    # room for one power-of-two padded segment per group bucket (csrc/nms.cu, kBySegs)
    h = ws.ensure(dev, 1, 1, max(2 * K + 512, 2))
    keep = torch.empty(max(K, 1), dtype=torch.int64, device=dev)
    n = torch.zeros(1, dtype=torch.int32, device=dev)
    lib = _lib.load()
    s = _stream(ltrb)
    with torch.cuda.device(dev):
        _lib.check(lib.yh_nms_by_scores(h, ltrb.data_ptr(), scores.data_ptr(), groups.data_ptr(), K,
                                        float(iou_threshold), keep.data_ptr(), n.data_ptr(), s))
    _, k = ws.read_counts(s, K_dev=n)
    return keep[:k]
