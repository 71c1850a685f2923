"""Pins oracle/evaluation.py against the reference's OWN known-answer tests (the only golden
vectors the reference holds on or next to the hot path):
  pred_gt_matching.rs:149-240   t_match_det_gt
  average_precision.rs:560-592  t_compute_by_detections   (AP == 1.0)
  average_precision.rs:594-639  t_mean_average_precision_cal  (0.9, 0.1, 0.19801980198019803)
  average_precision.rs:641-704  t_compute_by_prec_rec     (1.0 and 0.5908181818181819)
  utils/algorithm.rs:118-…      interpolate_test
The vectors below are the literals of those tests."""
import math

import pytest

from oracle import evaluation as ev

# pred_gt_matching.rs:152-161 / average_precision.rs:562-571 (gt: cls t l b r ; det: t l b r conf cls_conf cls)
GTS_A = """0 227.16200 219.68274 312.70200 410.39253
0 284.18624 189.21947 335.15290 404.17874
0 0.60445 237.66579 24.34890 415.77453
0 174.27155 155.53200 246.64890 359.78800
34 8.58000 330.53821 31.98000 411.12074"""
DETS_A = """175.30000 170.77000 245.34000 324.72000 0.99968 0.99998 0
284.07000 191.51000 336.73000 351.94000 0.98834 0.99999 0
229.29000 222.98000 314.37000 358.82000 0.98327 0.99990 0
0.35714 234.53000 29.80900 361.46000 0.89682 0.99831 0"""
# average_precision.rs:596-604
GTS_B = """39.00000 61.40888 27.67710 141.49845 230.31445
56.00000 0.22360 92.69645 58.11374 148.82400
56.00000 144.48242 43.56290 416.00021 231.43224
60.00000 0.00000 137.03310 412.75354 410.12421
40.00000 160.14066 101.55579 245.92610 240.79890"""
DETS_B = """159.15750 105.84630 247.27790 245.03130 0.99870 0.99960 40.00000
55.24000 31.11770 150.80330 362.72990 0.99670 0.99930 39.00000
200.69280 35.67050 411.24700 206.84590 0.78630 0.97070 56.00000"""


def parse_gts(text):
    rows = [[float(x) for x in line.split()] for line in text.splitlines()]
    return [int(r[0]) for r in rows], [r[1:5] for r in rows]


def parse_dets(text):
    rows = [[float(x) for x in line.split()] for line in text.splitlines()]
    return [r[0:4] for r in rows], [r[4] for r in rows], [r[5] for r in rows], [int(r[6]) for r in rows]


def iou_plus1(a, b):
    """the tests' own IoU (average_precision.rs:448-458): +1 pixel convention."""
    xa, ya, xb, yb = max(a[0], b[0]), max(a[1], b[1]), min(a[2], b[2]), min(a[3], b[3])
    inter = max(0.0, xb - xa + 1.0) * max(0.0, yb - ya + 1.0)
    aa = (a[2] - a[0] + 1.0) * (a[3] - a[1] + 1.0)
    ab = (b[2] - b[0] + 1.0) * (b[3] - b[1] + 1.0)
    return inter / (aa + ab - inter)


def match_d_g(dets, det_cls, gts, gt_cls):
    """average_precision.rs:460-494: class-filtered, strict '<' (first maximum)."""
    sel, ious = [], []
    for d, dc in zip(dets, det_cls):
        max_iou, g_sel = 0.0, 0
        for g, (gt, gc) in enumerate(zip(gts, gt_cls)):
            v = iou_plus1(d, gt)
            if gc == dc and max_iou < v:
                max_iou, g_sel = v, g
        sel.append(-1 if max_iou == 0.0 else g_sel)
        ious.append(0.0 if max_iou == 0.0 else iou_plus1(d, gts[g_sel]))
    return sel, ious


def test_t_match_det_gt():
    gt_cls, gts = parse_gts(GTS_A)
    dets, conf, cls_conf, det_cls = parse_dets(DETS_A)
    sel, ious = ev.match_det_gt(dets, gts)
    # the reference test's expectation: class-filtered max-IoU ground truth
    want = []
    for d, dc in zip(dets, det_cls):
        cands = [(ev.iou_with(d, g, ev.EPSILON), i) for i, (g, gc) in enumerate(zip(gts, gt_cls)) if gc == dc]
        best = max(cands, key=lambda t: (t[0], t[1]))
        want.append(-1 if abs(best[0]) <= ev.F64_EPS else best[1])
    assert sel == want == [3, 1, 0, 2]
    assert all(0.0 < v <= 1.0 for v in ious)


def test_t_compute_by_detections():
    gt_cls, gts = parse_gts(GTS_A)
    dets, conf, cls_conf, det_cls = parse_dets(DETS_A)
    sel, ious = match_d_g(dets, det_cls, gts, gt_cls)
    assert ev.ap_by_detections(sel, ious, cls_conf, 4, 0.5) == 1.0


def test_t_mean_average_precision_cal():
    gt_cls, gts = parse_gts(GTS_B)
    dets, conf, cls_conf, det_cls = parse_dets(DETS_B)
    sel, ious = match_d_g(dets, det_cls, gts, gt_cls)
    gt_cnt = {}
    for c in gt_cls:
        gt_cnt[c] = gt_cnt.get(c, 0) + 1
    classes = []
    for c in det_cls:
        if c not in classes:
            classes.append(c)
    per_class = {}
    for c in classes:
        idx = [i for i, dc in enumerate(det_cls) if dc == c]
        per_class[c] = ev.mean_ap([sel[i] for i in idx], [ious[i] for i in idx], [cls_conf[i] for i in idx],
                                  gt_cnt[c])
    total = sum(per_class.values()) / len(gt_cnt)
    assert abs(total - (0.9 + 0.1 + 0.19801980198019803) / len(gt_cnt)) <= ev.F64_EPS
    assert abs(per_class[40] - 0.9) < 1e-12 and abs(per_class[39] - 0.1) < 1e-12
    assert abs(per_class[56] - 0.19801980198019803) < 1e-12


def test_t_compute_by_prec_rec():
    assert ev.ap_by_prec_rec([(1.0, 1.0)], 101) == 1.0
    v = [(0.5, 0.625), (0.556, 0.625), (0.625, 0.625), (0.714, 0.625), (0.833, 0.625), (0.800, 0.500),
         (0.750, 0.375), (1.0, 0.375), (1.0, 0.250), (1.0, 0.125)]
    v.reverse()
    assert abs(ev.ap_by_prec_rec(v, 11) - 0.5908181818181819) <= ev.F64_EPS


def test_interpolate_test():
    out = ev.interpolate_stepwise_values([0.2], [(0.0, 3.0), (1.0, -5.0)])
    assert len(out) == 1 and out[0][0] == 0.2 and abs(out[0][1] + 5.0) <= 1e-5


def test_scalar_iou_known_cases():
    assert ev.iou_with([0, 0, 1, 1], [2, 2, 3, 3], ev.EPSILON) == 0.0          # disjoint
    assert ev.iou_with([0, 0, 1, 1], [1, 0, 2, 1], ev.EPSILON) == 0.0          # touching edge: b > t is strict
    assert math.isclose(ev.iou_with([0, 0, 2, 2], [1, 1, 3, 3], 0.0), 1.0 / 7.0)
    assert math.isclose(ev.iou_with([0, 0, 2, 2], [0, 0, 2, 2], ev.EPSILON), 4.0 / (4.0 + 1e-8))


def test_match_det_gt_last_maximum_and_zero_iou():
    gts = [[0, 0, 2, 2], [0, 0, 2, 2], [10, 10, 11, 11]]
    sel, ious = ev.match_det_gt([[0, 0, 2, 2], [20, 20, 21, 21]], gts)
    assert sel == [1, -1]            # max_by_key: the LAST of equal maxima; IoU 0 -> None
    assert ious[1] == 0.0
    with pytest.raises(ValueError):
        ev.match_det_gt([[0, 0, 1, 1]], [])
