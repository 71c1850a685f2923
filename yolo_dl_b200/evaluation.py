"""Evaluation metrics over the C ABI (yh_match_det_gt, yh_average_precision).

Mirrors yolo-dl/src/loss/pred_gt_matching.rs:52-142 (`MDetection`, `match_det_gt`) and
yolo-dl/src/loss/average_precision.rs:56-276 (`IntegralMethod`, `ApCalculator`,
`MeanApCalculator`).  Everything is f64 on the device, like the reference's R64 on the host.
"""
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import _lib


def _stream(t: torch.Tensor):
    return torch.cuda.current_stream(t.device).cuda_stream


def _i32(x, dev):
    return torch.as_tensor(x, dtype=torch.int32).to(dev).contiguous()


@dataclass
class DetGtMatch:
    """`Vec<DetBoxStruct>` (pred_gt_matching.rs:79-107) as two device tensors."""

    ground_truth: torch.Tensor  # [D] i32 row of the ground-truth tensor, -1 = None
    iou: torch.Tensor           # [D] f64 DetBoxStruct::iou


def match_det_gt(dets_tlbr: torch.Tensor, gts_tlbr: torch.Tensor,
                 det_offsets: Optional[Sequence[int]] = None,
                 gt_offsets: Optional[Sequence[int]] = None) -> DetGtMatch:
    """pred_gt_matching.rs:109-142.  dets_tlbr [D,4], gts_tlbr [G,4] f64 CUDA tensors of (t,l,b,r).
    Without offsets it is the reference call on one detection list and one ground-truth list; with
    [S+1] prefix offsets it runs S independent calls (images / classes) in one launch."""
    if not dets_tlbr.is_cuda:
        raise _lib.YoloHotError(_lib.YH_ERR_CUDA, "yolohot needs CUDA tensors; there is no CPU fallback")
    dev = dets_tlbr.device
    d = dets_tlbr.to(torch.float64).contiguous().view(-1, 4)
    g = gts_tlbr.to(device=dev, dtype=torch.float64).contiguous().view(-1, 4)
    D, G = d.shape[0], g.shape[0]
    doff = _i32(det_offsets if det_offsets is not None else [0, D], dev)
    goff = _i32(gt_offsets if gt_offsets is not None else [0, G], dev)
    if doff.numel() != goff.numel() or doff.numel() < 2:
        raise ValueError("det_offsets and gt_offsets must both have S+1 entries")
    out_gt = torch.full((D,), -1, dtype=torch.int32, device=dev)
    out_iou = torch.zeros(D, dtype=torch.float64, device=dev)
    lib = _lib.load()
    with torch.cuda.device(dev):
        _lib.check(lib.yh_match_det_gt(d.data_ptr() if D else None, doff.data_ptr(), D,
                                       g.data_ptr() if G else None, goff.data_ptr(), doff.numel() - 1,
                                       out_gt.data_ptr(), out_iou.data_ptr(), _stream(d)))
    return DetGtMatch(out_gt, out_iou)


@dataclass
class IntegralMethod:
    """average_precision.rs:56-60.  `Interpolation(n)`; `Continuous` is `todo!()` upstream."""

    n_points: Optional[int] = 101

    @staticmethod
    def interpolation(n: int) -> "IntegralMethod":
        return IntegralMethod(n)


class ApCalculator:
    def __init__(self, integral_method: IntegralMethod = None):
        m = integral_method or IntegralMethod(101)
        if m.n_points is None:
            raise NotImplementedError("IntegralMethod::Continuous is todo!() in the reference")
        if m.n_points < 1:
            raise ValueError(f"invalid number of interpolated points {m.n_points}")
        self.integral_method = m

    @staticmethod
    def new_coco() -> "ApCalculator":
        return ApCalculator(IntegralMethod(101))

    def compute_many(self, match: DetGtMatch, confidence: torch.Tensor, problem_offsets: Sequence[int],
                     num_ground_truth: Sequence[int], iou_thresholds: Sequence[float]) -> torch.Tensor:
        """[num_problems, num_thresholds] f64 APs in one launch (no synchronisation)."""
        dev = match.iou.device
        poff = [int(x) for x in problem_offsets]
        P = len(poff) - 1
        if P < 1 or len(num_ground_truth) != P:
            raise ValueError("problem_offsets needs P+1 entries and num_ground_truth P entries")
        dmax = max(poff[i + 1] - poff[i] for i in range(P))
        conf = confidence.to(device=dev, dtype=torch.float64).contiguous().view(-1)
        thr = torch.as_tensor(list(iou_thresholds), dtype=torch.float64).to(dev)
        out = torch.empty((P, thr.numel()), dtype=torch.float64, device=dev)
        poff_d, ngt_d = _i32(poff, dev), _i32(list(num_ground_truth), dev)
        lib = _lib.load()
        nscratch = int(lib.yh_average_precision_scratch_bytes(P, dmax))   # 0 up to 8192 detections
        scratch = torch.empty(nscratch, dtype=torch.uint8, device=dev) if nscratch else None
        with torch.cuda.device(dev):
            _lib.check(lib.yh_average_precision(
                match.ground_truth.data_ptr(), match.iou.data_ptr(), conf.data_ptr(), poff_d.data_ptr(),
                ngt_d.data_ptr(), P, dmax, thr.data_ptr(), thr.numel(), self.integral_method.n_points,
                out.data_ptr(), scratch.data_ptr() if nscratch else None, nscratch, _stream(conf)))
        return out

    def compute_by_detections(self, match: DetGtMatch, confidence: torch.Tensor, num_ground_truth: int,
                              iou_thresh: float) -> float:
        """average_precision.rs:157-219 for one detection list."""
        D = match.iou.shape[0]
        return float(self.compute_many(match, confidence, [0, D], [num_ground_truth], [iou_thresh])[0, 0])


class MeanApCalculator:
    def __init__(self, integral_method: IntegralMethod, iou_thresholds: Sequence[float]):
        if len(iou_thresholds) == 0:
            raise ValueError("iou_thresholds must be non-empty")
        self.ap_calculator = ApCalculator(integral_method)
        self.iou_thresholds: List[float] = [float(t) for t in iou_thresholds]

    @staticmethod
    def new_coco() -> "MeanApCalculator":
        return MeanApCalculator(IntegralMethod(101), [0.5 + i * 0.05 for i in range(10)])

    def compute_mean_ap(self, match: DetGtMatch, confidence: torch.Tensor, num_ground_truth: int) -> float:
        """average_precision.rs:252-275: the thresholds' APs summed in order, divided by their count."""
        D = match.iou.shape[0]
        aps = self.ap_calculator.compute_many(match, confidence, [0, D], [num_ground_truth],
                                              self.iou_thresholds)[0].tolist()
        s = 0.0
        for a in aps:
            s += a
        return s / len(aps)

    def compute_mean_ap_per_problem(self, match, confidence, problem_offsets, num_ground_truth) -> List[float]:
        aps = self.ap_calculator.compute_many(match, confidence, problem_offsets, num_ground_truth,
                                              self.iou_thresholds).tolist()
        out = []
        for row in aps:
            s = 0.0
            for a in row:
                s += a
            out.append(s / len(row))
        return out
