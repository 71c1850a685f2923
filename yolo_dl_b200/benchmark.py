"""YoloBenchmark over the C ABI (yh_benchmark).  Mirrors yolo-dl/src/loss/benchmark.rs:7-109.

The reference computes the objectness accuracy / recall / precision and the class accuracy of a
training step with ~12 ATen ops and five `i64::from(tensor)` device->host reads; here one kernel
counts, one 40-byte copy brings the five integers back and the ratios are formed in f64 on the
host with the reference's expressions (benchmark.rs:59-92)."""
from dataclasses import dataclass

import torch

from . import _lib
from .detect import MergedDenseDetection
from .tensors import MatchingOutput
from .workspace import Workspace


@dataclass
class YoloBenchmarkOutput:
    """benchmark.rs:103-109."""

    obj_accuracy: float
    obj_precision: float
    obj_recall: float
    class_accuracy: float


@dataclass
class YoloBenchmarkInit:
    """benchmark.rs:7-25."""

    iou_threshold: float
    confidence_threshold: float

    def build(self) -> "YoloBenchmark":
        return YoloBenchmark(float(self.iou_threshold), float(self.confidence_threshold))


def ratios_from_counts(all_count, all_pos, matched_count, matched_pos, class_hits) -> YoloBenchmarkOutput:
    """benchmark.rs:47-92 on the five integers."""
    all_neg = all_count - all_pos
    matched_neg = matched_count - matched_pos
    unmatched_neg = all_neg - matched_neg
    accuracy = (matched_pos + unmatched_neg) / all_count
    recall = matched_pos / matched_count if matched_count != 0 else 1.0
    precision = matched_pos / all_pos if all_pos != 0 else 1.0
    class_accuracy = class_hits / matched_count if matched_count != 0 else 1.0
    return YoloBenchmarkOutput(accuracy, precision, recall, class_accuracy)


class YoloBenchmark:
    def __init__(self, iou_threshold: float, confidence_threshold: float):
        self.iou_threshold = iou_threshold
        self.confidence_threshold = confidence_threshold
        self._lib = _lib.load()
        self._ws = Workspace()

    def counts(self, prediction: MergedDenseDetection, matchings: MatchingOutput) -> torch.Tensor:
        """device int64[5], no synchronisation."""
        dev = prediction.device()
        B, C = prediction.batch_size(), prediction.num_classes()
        rows = matchings.pred_rows
        if rows is None:  # a MatchingOutput not produced by YoloLoss: rebuild the contiguous rows
            p = matchings.pred
            rows = torch.cat([p.cycxhw.cy, p.cycxhw.cx, p.cycxhw.h, p.cycxhw.w, p.obj_logit, p.class_logit], 1)
        rows = rows.contiguous()
        M = rows.shape[0]
        cls = matchings.target.class_.reshape(-1).contiguous()
        m_dev = torch.tensor([M], dtype=torch.int32, device=dev)
        out = torch.empty(5, dtype=torch.int64, device=dev)
        self._ws.ensure(dev, B, prediction.num_instances())
        with torch.cuda.device(dev):
            _lib.check(self._lib.yh_benchmark(
                self._ws.handle, _lib.ptr_array(prediction.heads), _lib.make_layers(prediction.info),
                len(prediction.info), B, C, rows.data_ptr() if M else None, cls.data_ptr() if M else None,
                m_dev.data_ptr(), M, self.confidence_threshold, out.data_ptr(),
                torch.cuda.current_stream(dev).cuda_stream))
        self._keep = (rows, cls, m_dev)
        return out

    def forward(self, prediction: MergedDenseDetection, matchings: MatchingOutput, _inference=None) -> YoloBenchmarkOutput:
        c = self.counts(prediction, matchings).tolist()  # the one device->host read
        return ratios_from_counts(*c)
