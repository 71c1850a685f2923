"""CPU tests of the oracle itself: the reference's inline debug-assert restatements as property
tests, second implementations, and the committed golden vectors."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import box as obox
from oracle import detect as odet
from oracle import loss as oloss
from oracle import matcher as omatch
from oracle import nms as onms
from yolo_dl_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hotpath_small.npz")


def golden_inputs():
    g = np.load(GOLDEN)
    spec = synth.small_spec()
    heads = [torch.from_numpy(g[f"head{l}"]) for l in range(len(spec.shapes))]
    B = heads[0].shape[0]
    targets = [[] for _ in range(B)]
    for row in g["gt"]:
        targets[int(row[0])].append((row[1], row[2], row[3], row[4], int(row[5])))
    return g, spec, heads, targets


def test_golden_reproduced_by_oracle():
    g, spec, heads, targets = golden_inputs()
    infos = odet.make_infos(spec.shapes, spec.anchors)
    merged = odet.forward_merged(heads, infos, spec.num_classes)
    for k in ("cy", "cx", "h", "w"):
        np.testing.assert_allclose(merged[k].numpy(), g[f"dec_{k}"], rtol=1e-6, atol=1e-8)
    for metric in ("iou", "giou", "diou", "ciou"):
        r = oloss.yolo_loss(heads, infos, spec.num_classes, targets, oloss.LossConfig(box_metric=metric))
        got = np.array([float(r["losses"][k]) for k in ("total", "iou", "classification", "objectness")])
        np.testing.assert_allclose(got, g[f"{metric}_losses"], rtol=2e-6)
        assert np.array_equal(r["matches"][1], g["match_flats"])
        for l, gr in enumerate(r["grads"]):
            np.testing.assert_allclose(gr.numpy(), g[f"{metric}_grad{l}"], rtol=1e-5, atol=1e-9)
    for by_class, tag in ((False, "img"), (True, "cls")):
        k = onms.nms_forward(merged, 0.3, 0.5, by_class)
        assert np.array_equal(np.stack([k["batches"], k["classes"], k["instances"]], 1), g[f"nms_{tag}_rows"])


def test_decode_debug_assert_restatement():
    """detect_2d.rs:74-119."""
    spec = synth.small_spec()
    heads = synth.make_heads(spec, 2, 1)
    for h, (H, W), a in zip(heads, spec.shapes, spec.anchors):
        d = odet.detect2d_forward(h, spec.num_classes, a)
        v = h.view(2, -1, len(a), H, W)
        rows = torch.arange(H).view(1, 1, H, 1)
        cols = torch.arange(W).view(1, 1, 1, W)
        exp_cy = (v[:, 0].sigmoid() * 2 - 0.5 + rows) / H
        exp_cx = (v[:, 1].sigmoid() * 2 - 0.5 + cols) / W
        assert (d["cy"][:, 0] - exp_cy).abs().max() < 1e-6
        assert (d["cx"][:, 0] - exp_cx).abs().max() < 1e-6
        assert d["obj_logit"].data_ptr() == v[:, 4:5].data_ptr()


def test_merge_then_split_is_identity():
    """merged_dense_detection.rs:105-116."""
    spec = synth.small_spec()
    heads = synth.make_heads(spec, 2, 2)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    dets = [odet.detect2d_forward(h, spec.num_classes, i.anchors) for h, i in zip(heads, infos)]
    m = odet.merge(dets, spec.num_classes)
    for d, i in zip(dets, infos):
        sl = slice(i.flat_start, i.flat_start + i.size)
        assert torch.equal(m["cy"][:, :, sl].reshape(d["cy"].shape), d["cy"])
        assert torch.equal(m["class_logit"][:, :, sl].reshape(d["class_logit"].shape), d["class_logit"])
    # flat index formula (merged_dense_detection.rs:374-376)
    i = infos[1]
    f = odet.instance_to_flat(i, 1, 2, 3)
    assert float(m["cx"][0, 0, f]) == float(dets[1]["cx"][0, 0, 1, 2, 3])


def test_class_and_objectness_targets_debug_assert():
    """loss_.rs:342-368 (dense one-hot with pos=1-0.5*s) and loss_.rs:418-459 (scatter)."""
    spec = synth.small_spec()
    heads = synth.make_heads(spec, 2, 3)
    targets = synth.make_targets(spec, 2, 3, n_max=5)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    merged = odet.forward_merged(heads, infos, spec.num_classes)
    mb, mf, mg, mt, mc = omatch.match_bboxes(infos, targets)
    cfg = oloss.LossConfig(box_metric="giou", smooth_objectness_coef=0.4, smooth_classification_coef=0.1)
    losses, iou, pred, tdense, tobj = oloss.yolo_loss_from_matches(
        merged, torch.from_numpy(mb), torch.from_numpy(mf), torch.from_numpy(mt), torch.from_numpy(mc), cfg)
    pos = np.float32(1.0 - 0.5 * 0.1)
    neg = np.float32(1.0 - (1.0 - 0.5 * 0.1))
    exp = np.full(tdense.shape, neg, dtype=np.float32)
    exp[np.arange(len(mc)), mc] = pos
    assert np.array_equal(tdense.numpy(), exp)
    eo = np.zeros(tobj.shape, dtype=np.float32)
    sc = np.float32(1.0 - 0.4) + iou.detach().numpy().reshape(-1).clip(0, 1) * np.float32(0.4)
    eo[mb, 0, mf] = sc
    np.testing.assert_allclose(tobj.numpy(), eo, rtol=1e-6)


def test_matcher_python_vs_c_and_distance_bound():
    """two implementations agree; pred_target_matching.rs:161-174,219-235: matched cell within
    1 + snap_thresh cells of the GT centre."""
    for name, B, kw in (("small", 4, dict(n_max=12, clustered=0.5, adversarial=True)),
                        ("yolov4-608", 4, dict()),
                        ("yolov4-csp-640", 2, dict(n_min=300, n_max=300, clustered=0.3, adversarial=True))):
        spec = synth.SPECS[name]()
        infos = odet.make_infos(spec.shapes, spec.anchors)
        targets = synth.make_targets(spec, B, 17, **kw)
        for rect4 in (True, False):
            a = omatch.match_bboxes(infos, targets, rect4)
            b = omatch.match_bboxes_c(infos, targets, rect4)
            assert all(np.array_equal(x, y) for x, y in zip(a, b)), name
            mb, mf, mg, mt, mc = a
            assert len(set(zip(mb.tolist(), mf.tolist()))) == len(mb)  # one GT per slot
            for info in infos:
                sel = (mf >= info.flat_start) & (mf < info.flat_start + info.size)
                rem = mf[sel] - info.flat_start
                row = (rem % (info.H * info.W)) // info.W
                col = rem % info.W
                gcy = np.array([targets[b_][g][0] for b_, g in zip(mb[sel], mg[sel])])
                gcx = np.array([targets[b_][g][1] for b_, g in zip(mb[sel], mg[sel])])
                assert np.all(np.abs(gcy * info.H - row) <= 1.5)
                assert np.all(np.abs(gcx * info.W - col) <= 1.5)


def test_matcher_tie_keeps_first_gt_and_nearest_wins():
    infos = odet.make_infos([(4, 4)], [[(0.25, 0.25)]])
    # two GTs at the same centre: the first keeps the slot; a closer third GT takes it
    t = [[(0.55, 0.55, 0.2, 0.2, 0), (0.55, 0.55, 0.25, 0.25, 1)]]
    mb, mf, mg, mt, mc = omatch.match_bboxes(infos, t)
    assert set(mg.tolist()) == {0}
    t = [[(0.55, 0.55, 0.2, 0.2, 0), (0.62, 0.62, 0.25, 0.25, 1)]]
    mb, mf, mg, mt, mc = omatch.match_bboxes(infos, t)
    centre = odet.instance_to_flat(infos[0], 0, 2, 2)
    assert mg[list(mf).index(centre)] == 1  # (2.5,2.5)/4 = .625 is nearer to GT 1
    # anchor-ratio filter: a GT 5x the anchor is not matched at thresh 4
    assert len(omatch.match_bboxes(infos, [[(0.5, 0.5, 0.25 * 5, 0.25, 0)]])[0]) == 0
    # zero-sized boxes are ignored
    assert len(omatch.match_bboxes(infos, [[(0.5, 0.5, 0.0, 0.25, 0)]])[0]) == 0


def test_box_metrics_known_values_and_ciou_sign():
    def bx(cy, cx, h, w):
        return tuple(torch.tensor([[v]], dtype=torch.float32) for v in (cy, cx, h, w))
    a, b = bx(0.5, 0.5, 0.2, 0.2), bx(0.5, 0.6, 0.2, 0.2)
    iou = float(obox.iou_with(a, b))
    assert abs(iou - (0.02 / 0.06)) < 1e-6
    giou = float(obox.giou_with(a, b))
    assert abs(giou - (1 / 3 - (0.06 - 0.06) / 0.06)) < 1e-6
    diou = float(obox.diou_with(a, b))
    assert abs(diou - (1 / 3 - 0.01 / (0.04 + 0.09))) < 1e-6
    # CIoU = DIoU + alpha*v (cycxhw.rs:119), not minus
    c, d = bx(0.5, 0.5, 0.2, 0.4), bx(0.5, 0.5, 0.4, 0.2)
    v = 4 / math.pi ** 2 * (math.atan2(0.2, 0.4) - math.atan2(0.4, 0.2)) ** 2
    i2 = float(obox.iou_with(c, d))
    alpha = v / (1 - i2 + v)
    assert abs(float(obox.ciou_with(c, d)) - (float(obox.diou_with(c, d)) + alpha * v)) < 1e-6
    # disjoint boxes: IoU 0, eps keeps it finite
    assert float(obox.iou_with(bx(0.2, 0.2, 0.1, 0.1), bx(0.8, 0.8, 0.1, 0.1))) == 0.0


def test_bce_gradient_closed_form():
    """autograd of the restated loss == (sigmoid(x) - t) * w / numel on the objectness plane."""
    spec = synth.small_spec()
    heads = synth.make_heads(spec, 2, 5)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    r = oloss.yolo_loss(heads, infos, spec.num_classes, [[], []], oloss.LossConfig())
    N = sum(i.size for i in infos)
    for h, g, i in zip(heads, r["grads"], infos):
        x = h.view(2, -1, i.size)[:, 4]
        np.testing.assert_allclose(g.view(2, -1, i.size)[:, 4].numpy(), (torch.sigmoid(x) / (2 * N)).numpy(),
                                   rtol=1e-5)
        assert float(g.view(2, -1, i.size)[:, :4].abs().sum()) == 0.0


@pytest.mark.parametrize("K,groups", [(1, 1), (500, 1), (3000, 5), (3000, 80)])
def test_nms_oracle_vs_torchvision(K, groups):
    import torchvision
    ltrb, scores, grp = synth.make_nms_boxes(K, 3 + K, num_groups=groups, dup_frac=0.0)
    ref = torchvision.ops.batched_nms(torch.from_numpy(ltrb), torch.from_numpy(scores),
                                      torch.from_numpy(grp), 0.5).numpy()
    assert np.array_equal(onms.nms_by_scores(ltrb, scores, grp, 0.5), ref)


def test_nms_oracle_ties_keep_lower_index():
    ltrb = np.array([[0, 0, 1, 1]] * 3, dtype=np.float32)
    scores = np.array([0.7, 0.7, 0.7], dtype=np.float32)
    assert onms.nms_by_scores(ltrb, scores, np.zeros(3, np.int64), 0.5).tolist() == [0]
    assert onms.nms_by_scores(ltrb, scores, np.arange(3), 0.5).tolist() == [0, 1, 2]
    # suppress iff IoU > thr (not >=): two boxes with IoU exactly 0.5
    ltrb = np.array([[0, 0, 2, 1], [0, 0, 1, 1]], dtype=np.float32)
    assert onms.nms_by_scores(ltrb, np.array([0.9, 0.8], np.float32), np.zeros(2, np.int64), 0.5).tolist() == [0, 1]


def test_nms_forward_and_pick_on_yolov4_608():
    spec = synth.yolov4_608()
    heads = synth.make_heads(spec, 2, synth.SEED0)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    m = odet.forward_merged(heads, infos, spec.num_classes)
    a = onms.nms_forward(m, 0.1, 0.6, True)
    b = onms.nms_forward(m, 0.1, 0.6, True, use_torchvision=True)
    assert np.array_equal(a["instances"], b["instances"]) and np.array_equal(a["classes"], b["classes"])
    c = onms.nms_forward(m, 0.1, 0.6, True, compat_chunk65536=True)  # < 65536 candidates: same
    assert np.array_equal(a["instances"], c["instances"])
    p = onms.inference_pick(a)
    pairs = list(zip(p["batches"].tolist(), p["instances"].tolist()))
    assert len(pairs) == len(set(pairs)) and len(pairs) <= len(a["batches"])


def test_hausdorff_known_answers():
    """TLBRTensor::hausdorff_distance_to (tlbr.rs:137-177): hand-computed cases.  The reference's
    formula is the largest corner-to-corner displacement over the four corners, taken separately
    for the 'pred inside' and 'pred outside' signs."""
    from oracle import box as obox

    def t(*rows):
        a = torch.tensor(rows, dtype=torch.float32)
        return tuple(a[:, i:i + 1] for i in range(4))

    # identical boxes -> 0 ; same size shifted by (dy,dx)=(0.3,-0.4) -> 0.5
    p = t((0.5, 0.5, 0.2, 0.2), (0.5, 0.5, 0.2, 0.2), (0.5, 0.5, 0.2, 0.2))
    q = t((0.5, 0.5, 0.2, 0.2), (0.8, 0.1, 0.2, 0.2), (0.5, 0.5, 0.6, 0.4))
    d = obox.hausdorff_distance_to(p, q).view(-1)
    assert float(d[0]) == 0.0
    assert abs(float(d[1]) - 0.5) < 1e-6
    # concentric, target larger by 0.2 per side in y and 0.1 in x -> sqrt(0.2^2 + 0.1^2)
    assert abs(float(d[2]) - (0.2 ** 2 + 0.1 ** 2) ** 0.5) < 1e-6
    # independent numpy restatement on random boxes
    g = torch.Generator().manual_seed(5)
    a = torch.rand(200, 4, generator=g) * 0.5 + 0.1
    b = torch.rand(200, 4, generator=g) * 0.5 + 0.1
    got = obox.hausdorff_distance_to(tuple(a[:, i:i + 1] for i in range(4)),
                                     tuple(b[:, i:i + 1] for i in range(4))).view(-1).numpy()
    an, bn = a.numpy().astype(np.float64), b.numpy().astype(np.float64)

    def edges(x):
        return x[:, 0] - x[:, 2] / 2, x[:, 1] - x[:, 3] / 2, x[:, 0] + x[:, 2] / 2, x[:, 1] + x[:, 3] / 2
    pt, pl, pb, pr = edges(an)
    qt, ql, qb, qr = edges(bn)
    dd = [qt - pt, ql - pl, pb - qb, pr - qr]
    best = np.zeros(200)
    for sgn in (1.0, -1.0):
        c = [np.maximum(sgn * x, 0.0) for x in dd]
        for i, j in ((0, 1), (0, 3), (2, 1), (2, 3)):
            best = np.maximum(best, c[i] ** 2 + c[j] ** 2)
    assert np.allclose(got, np.sqrt(best), rtol=1e-5, atol=1e-6)


def test_loss_kinds_closed_forms():
    """Focal / L2 / CrossEntropy restatements (tch-modules focal_loss.rs:66-108, l2_loss.rs:13-15,
    cross_entropy.rs:19-58) against closed forms."""
    g = torch.Generator().manual_seed(11)
    x = torch.randn(6, 9, generator=g)
    t = torch.rand(6, 9, generator=g)
    bce = torch.nn.functional.binary_cross_entropy_with_logits(x, t, reduction="none")
    # gamma = 0 (YoloLossInit default): focal = bce * alpha_factor
    assert torch.allclose(oloss._focal(x, t, None, 0.0, "mean"), (bce * (t * 0.25 + (1 - t) * 0.75)).mean())
    p = torch.sigmoid(x)
    pt = t * p + (1 - t) * (1 - p)
    want = (bce * (t * 0.25 + (1 - t) * 0.75) * (1 - pt) ** 1.5).sum()
    assert torch.allclose(oloss._focal(x, t, None, 1.5, "sum"), want, rtol=1e-5)
    assert torch.allclose(oloss._l2(x, t, "mean"), ((x - t) ** 2).mean())
    assert torch.allclose(oloss._cross_entropy(x, t, "mean"),
                          torch.nn.functional.cross_entropy(x, t.argmax(1)))
    # empty input with mean reduction gives 0 (focal_loss.rs:79-81, cross_entropy.rs:45-47)
    e = torch.zeros(0, 9)
    assert float(oloss._focal(e, e, None, 1.5, "mean")) == 0.0
    assert float(oloss._cross_entropy(e, e, "mean")) == 0.0
