"""GPU parity: YoloLoss::forward + backward (yh_assign + yh_loss_fwd_bwd) against the oracle
(ATen op sequence + autograd on the CPU)."""
import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh
from oracle import loss as oloss
from yolo_dl_b200 import synth

from util import build, close, grad_close

pytestmark = pytest.mark.gpu


def _run(dev, spec, B, targets, seed=3, **kw):
    heads_cpu = synth.make_heads(spec, B, seed)
    heads, _, merged, infos = build(spec, heads_cpu, dev, requires_grad=True)
    init = yh.YoloLossInit(**kw)
    loss = init.build()
    out, aux = loss.forward(merged, targets)
    out.total_loss.backward()
    cfg = oloss.LossConfig(
        box_metric=init.box_metric, reduction=init.reduction, iou_loss_weight=init.iou_loss_weight,
        classification_loss_weight=init.classification_loss_weight,
        objectness_loss_weight=init.objectness_loss_weight,
        smooth_classification_coef=init.smooth_classification_coef,
        smooth_objectness_coef=init.smooth_objectness_coef,
        objectness_pos_weight=init.objectness_pos_weight,
        match_rect4=init.match_grid_method == "Rect4", anchor_scale_thresh=init.anchor_scale_thresh,
        focal_loss_gamma=init.focal_loss_gamma, objectness_loss_kind=init.objectness_loss_kind.lower(),
        classification_loss_kind=init.classification_loss_kind.lower())
    ref = oloss.yolo_loss(heads_cpu, infos, spec.num_classes, targets, cfg, use_c_matcher=True)
    return heads, out, aux, ref


def _compare(heads, out, aux, ref, what):
    L = ref["losses"]
    close(out.total_loss, L["total"], what=f"{what} total")
    close(out.iou_loss, L["iou"], what=f"{what} iou")
    close(out.classification_loss, L["classification"], what=f"{what} cls")
    close(out.objectness_loss, L["objectness"], what=f"{what} obj")
    assert np.array_equal(aux.matchings.pred_indexes.batches.cpu().numpy(), ref["matches"][0])
    assert np.array_equal(aux.matchings.pred_indexes.flats.cpu().numpy(), ref["matches"][1])
    # box metric per matched row: |a-b| <= 1e-5 (scores live in [-2,2]; relative error is
    # unbounded at the zero crossings of GIoU/DIoU/CIoU)
    if aux.iou_score is not None:
        close(aux.iou_score, ref["iou_score"], rtol=1e-5, atol=1e-5, what=f"{what} iou_score")
    p = aux.matchings.pred
    close(p.cycxhw.cy, ref["pred"]["cy"], what="pred cy")
    close(p.cycxhw.w, ref["pred"]["w"], what="pred w")
    assert torch.equal(p.class_logit.cpu(), ref["pred"]["class_logit"])
    for l, (h, g) in enumerate(zip(heads, ref["grads"])):
        grad_close(h.grad, g, what=f"{what} grad layer {l}")


@pytest.mark.parametrize("metric", ["iou", "giou", "diou", "ciou", "hausdorff"])
def test_loss_small_all_metrics(dev, metric):
    spec = synth.small_spec()
    targets = synth.make_targets(spec, 4, 31, n_max=8, clustered=0.3, adversarial=True)
    heads, out, aux, ref = _run(dev, spec, 4, targets, box_metric=metric)
    _compare(heads, out, aux, ref, metric)


@pytest.mark.parametrize("metric", ["ciou", "diou"])
def test_loss_yolov4_608_batch8(dev, metric):
    """BASELINE config 1 per-GPU share: YOLOv4-608, batch 8, CIoU and the default DIoU."""
    spec = synth.yolov4_608()
    targets = synth.make_targets(spec, 8, synth.SEED0 + 1)
    heads, out, aux, ref = _run(dev, spec, 8, targets, seed=synth.SEED0 + 1, box_metric=metric)
    _compare(heads, out, aux, ref, f"608 {metric}")


def test_loss_hausdorff_yolov4_608_and_no_score(dev):
    """BoxMetric::Hausdorff (loss_.rs:285-288): the distance is the loss, the auxiliary score is
    None and objectness smoothing falls back to the constant 1 - s_o."""
    spec = synth.yolov4_608()
    targets = synth.make_targets(spec, 4, 19, n_max=30)
    heads, out, aux, ref = _run(dev, spec, 4, targets, box_metric="hausdorff",
                                smooth_objectness_coef=0.2)
    assert aux.iou_score is None
    _compare(heads, out, aux, ref, "hausdorff 608")


@pytest.mark.parametrize("kw", [
    dict(objectness_loss_kind="Focal", classification_loss_kind="Focal"),                       # gamma 0 default
    dict(objectness_loss_kind="Focal", classification_loss_kind="Focal", focal_loss_gamma=1.5),
    dict(objectness_loss_kind="Focal", focal_loss_gamma=2.0, objectness_pos_weight=1.7,
         smooth_objectness_coef=0.4, box_metric="ciou"),
    dict(objectness_loss_kind="L2", classification_loss_kind="L2"),
    dict(classification_loss_kind="CrossEntropy"),
    dict(classification_loss_kind="CrossEntropy", reduction="sum", objectness_loss_kind="L2"),
    dict(classification_loss_kind="CrossEntropy", smooth_classification_coef=1.0),            # pos == neg: arg-max is 0
], ids=lambda kw: ",".join(f"{k}={v}" for k, v in kw.items()))
def test_loss_kinds_small(dev, kw):
    """ObjectnessLossKind / ClassificationLossKind (loss_.rs:89-145): Focal, L2, CrossEntropy."""
    spec = synth.small_spec()
    targets = synth.make_targets(spec, 3, 43, n_max=7, clustered=0.3)
    heads, out, aux, ref = _run(dev, spec, 3, targets, **kw)
    _compare(heads, out, aux, ref, str(kw))


def test_loss_kinds_yolov4_608(dev):
    spec = synth.yolov4_608()
    targets = synth.make_targets(spec, 4, 23)
    heads, out, aux, ref = _run(dev, spec, 4, targets, box_metric="ciou", objectness_loss_kind="Focal",
                                classification_loss_kind="CrossEntropy", focal_loss_gamma=1.5)
    _compare(heads, out, aux, ref, "608 focal+ce")


def test_loss_options_sum_smoothing_posweight(dev):
    spec = synth.small_spec()
    targets = synth.make_targets(spec, 3, 41, n_max=6)
    for kw in (dict(reduction="sum"),
               dict(smooth_objectness_coef=0.7, smooth_classification_coef=0.2, box_metric="giou"),
               dict(objectness_pos_weight=2.5, box_metric="ciou", smooth_objectness_coef=0.3),
               dict(match_grid_method="Rect2", anchor_scale_thresh=2.0, iou_loss_weight=1.0,
                    classification_loss_weight=0.0)):
        heads, out, aux, ref = _run(dev, spec, 3, targets, **kw)
        _compare(heads, out, aux, ref, str(kw))


def test_loss_no_targets(dev):
    """M = 0: iou and class loss are 0 (loss_.rs:309-314, bce_with_logits_loss.rs:74-77), the
    objectness loss and its gradient remain."""
    spec = synth.small_spec()
    heads, out, aux, ref = _run(dev, spec, 2, [[], []])
    assert aux.matchings.num_samples() == 0
    assert float(out.iou_loss) == 0.0 and float(out.classification_loss) == 0.0
    _compare(heads, out, aux, ref, "empty")


def test_loss_forward_only_matches_and_leaves_no_grad(dev):
    spec = synth.small_spec()
    targets = synth.make_targets(spec, 2, 51, n_max=5)
    heads_cpu = synth.make_heads(spec, 2, 3)
    heads, _, merged, infos = build(spec, heads_cpu, dev, requires_grad=False)
    out, aux = yh.YoloLossInit().build().forward(merged, targets)
    ref = oloss.yolo_loss(heads_cpu, infos, spec.num_classes, targets, oloss.LossConfig(), want_grad=False)
    close(out.total_loss, ref["losses"]["total"], what="fwd-only total")
    assert not out.total_loss.requires_grad


def test_loss_deterministic_bits(dev):
    """fixed-order reductions: two runs give identical bits (losses and gradients)."""
    spec = synth.yolov4_608()
    targets = synth.make_targets(spec, 4, 61)
    h1, o1, a1, _ = None, None, None, None
    outs = []
    for _ in range(2):
        heads_cpu = synth.make_heads(spec, 4, 9)
        heads, _, merged, _ = build(spec, heads_cpu, dev, requires_grad=True)
        out, aux = yh.YoloLossInit(box_metric="ciou").build().forward(merged, targets)
        out.total_loss.backward()
        outs.append((out, [h.grad.clone() for h in heads]))
    assert torch.equal(outs[0][0].total_loss, outs[1][0].total_loss)
    for a, b in zip(outs[0][1], outs[1][1]):
        assert torch.equal(a, b)


def test_loss_validation_messages():
    with pytest.raises(ValueError, match=r"smooth_classification_coef must be in range \[0, 1\]"):
        yh.YoloLossInit(smooth_classification_coef=1.5).build()
    with pytest.raises(ValueError, match="iou_loss_weight must be non-negative"):
        yh.YoloLossInit(iou_loss_weight=-1.0).build()


def test_full_size_properties_batch64(dev):
    """BASELINE size (YOLOv4-608, batch 64) through size-independent properties: the class target
    row sums (loss_.rs:342-368), the objectness gradient identity d/dx = (sigmoid(x)-t)*w/(B*N),
    zero gradient outside matched rows / the objectness plane, linearity of the gradient in the
    loss weights."""
    spec = synth.yolov4_608()
    B = 64
    heads = synth.make_heads(spec, B, 77, device=dev)
    targets = synth.make_targets(spec, B, 78)
    detects = [yh.Detect2DInit(spec.num_classes, a).build() for a in spec.anchors]
    for h in heads:
        h.requires_grad_(True)
    merged = yh.MergedDenseDetection.from_heads(heads, detects)
    loss = yh.YoloLossInit(box_metric="ciou").build()
    out, aux = loss.forward(merged, targets)
    out.total_loss.backward()
    N = merged.num_instances()
    M = aux.matchings.num_samples()
    assert M > 0
    bt, fl = aux.matchings.pred_indexes.batches, aux.matchings.pred_indexes.flats
    tgt = torch.zeros((B, N), device=dev)
    tgt[bt, fl] = 1.0
    E = spec.num_classes + 5
    total_nonobj = 0
    for h, info in zip(heads, merged.info):
        g = h.grad.view(B, E, info.size)
        x = h.detach().view(B, E, info.size)[:, 4]
        t = tgt[:, info.flat_start:info.flat_start + info.size]
        expect = (torch.sigmoid(x) - t) / (B * N)
        assert torch.allclose(g[:, 4], expect, rtol=1e-5, atol=1e-12)
        other = torch.cat([g[:, :4], g[:, 5:]], 1)
        total_nonobj += int((other != 0).any(1).sum())
        # gradient outside matched instances is exactly zero
        assert bool((other.abs().sum(1)[t == 0] == 0).all())
    assert total_nonobj <= M
    # linearity: doubling every weight doubles the gradient bit-exactly (powers of two)
    g1 = [h.grad.clone() for h in heads]
    for h in heads:
        h.grad = None
    loss2 = yh.YoloLossInit(box_metric="ciou", iou_loss_weight=0.1, classification_loss_weight=1.16,
                            objectness_loss_weight=2.0).build()
    out2, _ = loss2.forward(merged, targets)
    out2.total_loss.backward()
    for a, h in zip(g1, heads):
        assert torch.equal(a * 2, h.grad)
    assert torch.allclose(out2.total_loss, out.total_loss * 2, rtol=1e-6)


def test_unsorted_match_list_is_reported(dev):
    """The gradient scatter merges neighbouring matched cells and therefore needs the match list
    in yh_assign's order (ascending (batch, flat)); any other order must fail loudly at the next
    yh_read_counts, not scatter wrong values quietly."""
    from yolo_dl_b200 import _lib
    from yolo_dl_b200.pipeline import DeviceStep
    spec = synth.small_spec()
    B = 3
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 5)]
    gt_h, off_h = yh.pack_labels(synth.make_targets(spec, B, 5, n_max=6))
    step = DeviceStep(spec, B, dev)
    s = torch.cuda.current_stream(dev).cuda_stream
    step._assign(gt_h.to(dev), off_h.to(dev), s)
    hp = _lib.ptr_array(heads)
    step._loss(step.lib.yh_loss_fwd_bwd, hp, s)
    M, _ = step.counts()  # sorted list: fine
    assert M > 2
    for t in (step.m_batch, step.m_flat, step.m_class):
        t[:M] = t[:M].flip(0)
    step.m_target[:M] = step.m_target[:M].flip(0)
    step._loss(step.lib.yh_loss_fwd_bwd, hp, s)
    with pytest.raises(yh.YoloHotError) as e:
        step.counts()
    assert "not sorted" in str(e.value)
    step._assign(gt_h.to(dev), off_h.to(dev), s)
    step._loss(step.lib.yh_loss_fwd_bwd, hp, s)
    assert step.counts()[0] == M  # the flag is cleared by the read
