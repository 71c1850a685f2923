"""yh_match_det_gt / yh_average_precision against the reference's known-answer tests and, on
random problems, against oracle/evaluation.py (which those same tests pin).  f64 throughout:
matched indices and AP values are compared for EXACT equality."""
import numpy as np
import pytest
import torch

from oracle import evaluation as ev
from yolo_dl_b200 import evaluation as yev
from test_oracle_eval import DETS_A, DETS_B, GTS_A, GTS_B, match_d_g, parse_dets, parse_gts

pytestmark = pytest.mark.gpu


def dmatch(dev, sel, ious):
    return yev.DetGtMatch(torch.tensor(sel, dtype=torch.int32, device=dev),
                          torch.tensor(ious, dtype=torch.float64, device=dev))


def test_reference_t_match_det_gt(dev):
    gt_cls, gts = parse_gts(GTS_A)
    dets, conf, cls_conf, det_cls = parse_dets(DETS_A)
    m = yev.match_det_gt(torch.tensor(dets, dtype=torch.float64, device=dev),
                         torch.tensor(gts, dtype=torch.float64, device=dev))
    sel, ious = ev.match_det_gt(dets, gts)
    assert m.ground_truth.tolist() == sel == [3, 1, 0, 2]
    assert m.iou.tolist() == ious


def test_reference_t_compute_by_detections(dev):
    gt_cls, gts = parse_gts(GTS_A)
    dets, conf, cls_conf, det_cls = parse_dets(DETS_A)
    sel, ious = match_d_g(dets, det_cls, gts, gt_cls)
    ap = yev.ApCalculator.new_coco().compute_by_detections(
        dmatch(dev, sel, ious), torch.tensor(cls_conf, dtype=torch.float64, device=dev), 4, 0.5)
    assert ap == 1.0


def test_reference_t_mean_average_precision_cal(dev):
    gt_cls, gts = parse_gts(GTS_B)
    dets, conf, cls_conf, det_cls = parse_dets(DETS_B)
    sel, ious = match_d_g(dets, det_cls, gts, gt_cls)
    gt_cnt = {}
    for c in gt_cls:
        gt_cnt[c] = gt_cnt.get(c, 0) + 1
    # one problem per detected class, detections grouped by class
    classes = sorted(set(det_cls))
    order = [i for c in classes for i, dc in enumerate(det_cls) if dc == c]
    offs = [0]
    for c in classes:
        offs.append(offs[-1] + sum(1 for dc in det_cls if dc == c))
    m = dmatch(dev, [sel[i] for i in order], [ious[i] for i in order])
    cf = torch.tensor([cls_conf[i] for i in order], dtype=torch.float64, device=dev)
    per = yev.MeanApCalculator.new_coco().compute_mean_ap_per_problem(m, cf, offs, [gt_cnt[c] for c in classes])
    got = dict(zip(classes, per))
    assert abs(got[40] - 0.9) < 1e-12 and abs(got[39] - 0.1) < 1e-12
    assert abs(got[56] - 0.19801980198019803) < 1e-12
    total = sum(per) / len(gt_cnt)
    assert abs(total - (0.9 + 0.1 + 0.19801980198019803) / len(gt_cnt)) <= 1e-15
    # and exactly the oracle's bits
    for c in classes:
        idx = [i for i, dc in enumerate(det_cls) if dc == c]
        assert got[c] == ev.mean_ap([sel[i] for i in idx], [ious[i] for i in idx], [cls_conf[i] for i in idx],
                                    gt_cnt[c])


def random_boxes(rng, n, scale=400.0):
    t = rng.uniform(0, scale, n)
    l = rng.uniform(0, scale, n)
    h = rng.uniform(5, 120, n)
    w = rng.uniform(5, 120, n)
    return np.stack([t, l, t + h, l + w], 1)


@pytest.mark.parametrize("seed,S,dmax,gmax", [(0, 1, 40, 9), (1, 7, 300, 40), (2, 64, 60, 12), (3, 3, 2000, 70)])
def test_match_det_gt_random_segments(dev, seed, S, dmax, gmax):
    rng = np.random.default_rng(seed)
    dets, gts, doff, goff = [], [], [0], [0]
    for s in range(S):
        nd, ng = int(rng.integers(0, dmax + 1)), int(rng.integers(1, gmax + 1))
        g = random_boxes(rng, ng)
        d = random_boxes(rng, nd)
        k = min(nd, ng)
        if k:  # some detections near / exactly on ground truths, some duplicated ground truths (ties)
            d[:k] = g[:k] + rng.normal(0, 3.0, (k, 4))
            d[0] = g[0]
        if ng >= 2:
            g[-1] = g[0]
        dets.append(d)
        gts.append(g)
        doff.append(doff[-1] + nd)
        goff.append(goff[-1] + ng)
    dets, gts = np.concatenate(dets), np.concatenate(gts)
    m = yev.match_det_gt(torch.from_numpy(dets).to(dev), torch.from_numpy(gts).to(dev), doff, goff)
    got_sel, got_iou = m.ground_truth.tolist(), m.iou.tolist()
    for s in range(S):
        sel, ious = ev.match_det_gt(dets[doff[s]:doff[s + 1]].tolist(), gts[goff[s]:goff[s + 1]].tolist())
        sel = [-1 if x < 0 else x + goff[s] for x in sel]
        assert got_sel[doff[s]:doff[s + 1]] == sel
        assert got_iou[doff[s]:doff[s + 1]] == ious


def test_match_det_gt_segment_without_ground_truth(dev):
    d = torch.tensor([[0.0, 0.0, 1.0, 1.0]], dtype=torch.float64, device=dev)
    g = torch.zeros((0, 4), dtype=torch.float64, device=dev)
    m = yev.match_det_gt(d, g)
    assert m.ground_truth.tolist() == [-1] and m.iou.tolist() == [0.0]


@pytest.mark.parametrize("seed,P,dmax", [(0, 1, 50), (1, 12, 400), (2, 80, 90), (3, 2, 8192), (4, 5, 1),
                                         (5, 3, 8193), (6, 2, 30000)])
def test_average_precision_random_equals_oracle(dev, seed, P, dmax):
    """seeds 5, 6: more than 8192 detections in one problem (a COCO-val "person" class): the
    per-detection arrays move from shared memory to the caller's scratch, same bits out."""
    rng = np.random.default_rng(100 + seed)
    sel, iou, conf, offs, ngt = [], [], [], [0], []
    for p in range(P):
        nd = dmax if (seed in (3, 5, 6) and p == 0) else int(rng.integers(1, dmax + 1))
        g = int(rng.integers(1, 40))
        s = rng.integers(-1, g, nd)
        v = np.where(s < 0, 0.0, rng.uniform(0.05, 1.0, nd))
        v = np.round(v, 2)                      # many equal IoUs, some exactly on a threshold
        c = np.round(rng.uniform(0.0, 1.0, nd), 2)  # many equal confidences
        sel += s.tolist(); iou += v.tolist(); conf += c.tolist()
        offs.append(offs[-1] + nd)
        ngt.append(max(g, int(rng.integers(g, g + 5))))
    thr = ev.coco_iou_thresholds()
    calc = yev.ApCalculator.new_coco()
    got = calc.compute_many(dmatch(dev, sel, iou), torch.tensor(conf, dtype=torch.float64, device=dev),
                            offs, ngt, thr).tolist()
    for p in range(P):
        a, b = offs[p], offs[p + 1]
        local = [x for x in sel[a:b]]
        for k, t in enumerate(thr):
            want = ev.ap_by_detections(local, iou[a:b], conf[a:b], ngt[p], t)
            assert got[p][k] == want, (p, k, got[p][k], want)


def test_average_precision_other_point_counts_and_full_recall(dev):
    # every detection a true positive with recall reaching exactly 1.0 (the `latter.x == 1.0` arm)
    sel, iou, conf = [0, 1, 2, 3], [0.9, 0.8, 0.7, 0.6], [0.9, 0.8, 0.7, 0.6]
    for n in (2, 11, 101, 1000):
        calc = yev.ApCalculator(yev.IntegralMethod(n))
        got = calc.compute_by_detections(dmatch(dev, sel, iou), torch.tensor(conf, device=dev), 4, 0.5)
        assert got == ev.ap_by_detections(sel, iou, conf, 4, 0.5, n)
    assert yev.ApCalculator.new_coco().compute_by_detections(
        dmatch(dev, sel, iou), torch.tensor(conf, device=dev), 4, 0.5) == 1.0


def test_evaluation_argument_validation(dev):
    with pytest.raises(ValueError):
        yev.ApCalculator(yev.IntegralMethod(0))
    with pytest.raises(ValueError):
        yev.MeanApCalculator(yev.IntegralMethod(101), [])
    # more than 8192 detections without scratch: refused with the size to pass, not computed wrongly
    from yolo_dl_b200 import _lib
    lib = _lib.load()
    assert lib.yh_average_precision_scratch_bytes(4, 8192) == 0
    need = lib.yh_average_precision_scratch_bytes(1, 9000)
    assert need >= 9000 * 17
    m = dmatch(dev, [0] * 9000, [0.5] * 9000)
    cf = torch.zeros(9000, dtype=torch.float64, device=dev)
    off = torch.tensor([0, 9000], dtype=torch.int32, device=dev)
    ng = torch.tensor([3], dtype=torch.int32, device=dev)
    th = torch.tensor([0.5], dtype=torch.float64, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    rc = lib.yh_average_precision(m.ground_truth.data_ptr(), m.iou.data_ptr(), cf.data_ptr(), off.data_ptr(),
                                  ng.data_ptr(), 1, 9000, th.data_ptr(), 1, 101, out.data_ptr(), None, 0, None)
    assert rc == _lib.YH_ERR_CAPACITY and b"scratch" in lib.yh_last_error()
