"""Evaluation metrics restated in plain Python floats (IEEE f64 like the reference's R64):
  bbox::RectFloat::iou_with          bbox/src/rect.rs:113-140
  match_det_gt                       yolo-dl/src/loss/pred_gt_matching.rs:109-142
  ApCalculator / MeanApCalculator    yolo-dl/src/loss/average_precision.rs:63-276
  interpolate_stepwise_values        yolo-dl/src/utils/algorithm.rs:38-91
Test infrastructure only (see oracle/__init__.py).  PINNED by the reference's own known-answer
tests (pred_gt_matching.rs:149-240, average_precision.rs:560-705): tests/test_oracle_eval.py.
"""
from typing import List, Optional, Sequence, Tuple

EPSILON = 1e-8          # pred_gt_matching.rs:5
F64_EPS = 2.220446049250313e-16  # approx::abs_diff_eq! default epsilon for f64


def iou_with(a: Sequence[float], b: Sequence[float], epsilon: float) -> float:
    """a, b = (t, l, b, r).  rect.rs:113-140."""
    t, l = max(a[0], b[0]), max(a[1], b[1])
    bb, r = min(a[2], b[2]), min(a[3], b[3])
    inter = (bb - t) * (r - l) if (bb > t and r > l) else 0.0
    area_a = (a[2] - a[0]) * (a[3] - a[1])
    area_b = (b[2] - b[0]) * (b[3] - b[1])
    union = area_a + area_b - inter + epsilon
    return inter / union


def match_det_gt(dets_tlbr: Sequence[Sequence[float]], gts_tlbr: Sequence[Sequence[float]]
                 ) -> Tuple[List[int], List[float]]:
    """For every detection the ground truth of maximum IoU over ALL ground truths (no class filter,
    pred_gt_matching.rs:120-128); Iterator::max_by_key keeps the LAST maximum; the match is dropped
    (None, here -1) when that IoU is 0 within f64 epsilon (:131).  Returns (gt index | -1, IoU with
    the matched gt | 0.0) — the `DetBoxStruct.ground_truth` / `DetectionForAp::iou` pair."""
    sel, ious = [], []
    for d in dets_tlbr:
        best, best_iou = -1, None
        for g, gt in enumerate(gts_tlbr):
            v = iou_with(d, gt, EPSILON)
            if best_iou is None or v >= best_iou:
                best, best_iou = g, v
        if best_iou is None:
            raise ValueError("match_det_gt needs at least one ground truth (unwrap() on empty, :128)")
        if abs(best_iou - 0.0) <= F64_EPS:
            sel.append(-1)
            ious.append(0.0)
        else:
            sel.append(best)
            # DetBoxStruct::iou (:96-106) recomputes gt.iou_with(det): same value, operands swapped
            ious.append(iou_with(gts_tlbr[best], d, EPSILON))
    return sel, ious


def interpolate_stepwise_values(points: Sequence[float], values: Sequence[Tuple[float, float]]):
    """algorithm.rs:38-91, values = [(x, y)] ordered by x."""
    pts = iter(points)
    pairs = iter(zip(values, values[1:]))
    former, latter = next(pairs)
    inter_x = next(pts)
    out = []
    while True:
        c = (former[0] <= inter_x, latter[0] <= inter_x)
        if c == (False, False):
            nxt = next(pts, None)
            if nxt is not None:
                inter_x = nxt
            else:
                if abs(inter_x - former[0]) <= F64_EPS:
                    out.append((former[0], former[1]))
                break
        elif c == (True, False):
            out.append((inter_x, latter[1]))
            nxt = next(pts, None)
            if nxt is not None:
                inter_x = nxt
            else:
                break
        elif c == (True, True):
            nxt = next(pairs, None)
            if nxt is not None:
                former, latter = nxt
            else:
                if latter[0] == 1.0:
                    out.append((latter[0], latter[1]))
                elif abs(inter_x - latter[0]) <= F64_EPS:
                    out.append((latter[0], 0.0))
                break
        else:
            raise AssertionError("unreachable (algorithm.rs:86)")
    return out


def ap_by_prec_rec(sorted_prec_rec: Sequence[Tuple[float, float]], n_points: int = 101) -> float:
    """ApCalculator::compute_by_prec_rec (average_precision.rs:87-155); input (precision, recall)
    ordered by non-decreasing recall."""
    max_recall = sorted_prec_rec[-1][1]
    first = (0.0, 0.0)
    last = (0.0, min(max_recall + 1e-3, 1.0))
    seq = [first] + list(sorted_prec_rec) + [last]
    lst: List[List[float]] = []  # [precision, recall]
    for precision, recall in reversed(seq):
        if lst:
            l = lst[-1]
            max_precision = max(l[0], precision)
            if l[1] == recall:
                l[0] = max(l[0], precision)
            else:
                lst.append([max_precision, recall])
        else:
            lst.append([precision, recall])
    lst.reverse()
    enveloped = [(r, p) for p, r in lst]  # AsXY: x = recall, y = precision
    points = [i / (n_points - 1) for i in range(n_points)] if n_points > 1 else [float("nan")]
    inter = interpolate_stepwise_values(points, enveloped)
    s = 0.0
    for _, p in inter:
        s += p
    return s / n_points


def ap_by_detections(gt_index: Sequence[int], iou: Sequence[float], confidence: Sequence[float],
                     num_ground_truth: int, iou_thresh: float, n_points: int = 101) -> float:
    """ApCalculator::compute_by_detections (average_precision.rs:157-219).  Detection i is matched
    to ground truth gt_index[i] (-1 = None) with IoU iou[i].  Per ground truth (None is a group of
    its own, :171-174) the detection of highest IoU is a true positive iff IoU >= thresh, the rest
    false positives; then a stable sort by decreasing confidence and the running precision/recall.
    Group iteration order of the reference is a HashMap's (unspecified); it only matters among
    equal confidences, where this restatement uses ascending group key, None first."""
    groups = {}
    for i, g in enumerate(gt_index):
        groups.setdefault(g, []).append(i)
    flagged = []
    for g in sorted(groups):
        members = sorted(groups[g], key=lambda i: -iou[i])  # stable, like sort_by_cached_key
        for rank, i in enumerate(members):
            flagged.append((i, rank == 0 and iou[i] >= iou_thresh))
    flagged.sort(key=lambda t: -confidence[t[0]])  # stable
    tp = fp = 0
    prec_rec = []
    for _, is_tp in flagged:
        if is_tp:
            tp += 1
        else:
            fp += 1
        prec_rec.append((tp / (tp + fp), tp / num_ground_truth))
    return ap_by_prec_rec(prec_rec, n_points)


def coco_iou_thresholds() -> List[float]:
    return [0.5 + i * 0.05 for i in range(10)]  # MeanApCalculator::new_coco (:229-238)


def mean_ap(gt_index, iou, confidence, num_ground_truth: int, iou_thresholds: Optional[Sequence[float]] = None,
            n_points: int = 101) -> float:
    """MeanApCalculator::compute_mean_ap (:252-275)."""
    th = list(iou_thresholds) if iou_thresholds is not None else coco_iou_thresholds()
    s = 0.0
    for t in th:
        s += ap_by_detections(gt_index, iou, confidence, num_ground_truth, t, n_points)
    return s / len(th)
