"""YoloLoss::forward restated with the reference's ATen op order
(yolo-dl/src/loss/loss_.rs:227-468; tch-modules/src/bce_with_logits_loss.rs:60-85).

Test infrastructure only (see oracle/__init__.py).
"""
from dataclasses import dataclass
from typing import Optional, Sequence

import torch
import torch.nn.functional as F

from . import box as obox
from . import detect as odet
from . import matcher as omatch


@dataclass
class LossConfig:
    """YoloLossInit defaults (loss_.rs:169-189)."""

    box_metric: str = "diou"
    reduction: str = "mean"
    iou_loss_weight: float = 0.05
    classification_loss_weight: float = 0.58
    objectness_loss_weight: float = 1.0
    smooth_classification_coef: float = 0.01
    smooth_objectness_coef: float = 0.0
    objectness_pos_weight: Optional[float] = None
    match_rect4: bool = True
    anchor_scale_thresh: float = 4.0
    focal_loss_gamma: float = 0.0             # loss_.rs:176
    objectness_loss_kind: str = "bce"         # bce | focal | l2            (loss_.rs:115-145)
    classification_loss_kind: str = "bce"     # bce | focal | l2 | crossentropy (loss_.rs:89-113)


def _bce(inp, target, pos_weight, reduction):
    """BceWithLogitsLoss::forward (bce_with_logits_loss.rs:60-85)."""
    if inp.numel() == 0 and reduction == "mean":
        return torch.zeros((), dtype=torch.float32, device=inp.device)
    return F.binary_cross_entropy_with_logits(inp, target, weight=None, pos_weight=pos_weight,
                                              reduction=reduction)


def _reduce(loss, reduction):
    return loss.mean() if reduction == "mean" else loss.sum()


def _focal(inp, target, pos_weight, gamma, reduction, alpha=0.25):
    """FocalLoss::forward (tch-modules/src/focal_loss.rs:66-108) around BceWithLogitsLoss with
    Reduction::None (loss_.rs:97-111,127-143); alpha = FocalLossInit::default's 0.25."""
    if inp.numel() == 0 and reduction == "mean":
        return torch.zeros((), dtype=torch.float32, device=inp.device)
    orig_loss = _bce(inp, target, pos_weight, "none")
    input_prob = inp.sigmoid()
    p_t = target * input_prob + (1.0 - target) * (1.0 - input_prob)
    alpha_factor = target * alpha + (1.0 - target) * (1.0 - alpha)
    modulating_factor = (-p_t + 1.0).pow(torch.tensor(gamma, dtype=torch.float64))
    loss = orig_loss * alpha_factor * modulating_factor
    return loss if reduction == "none" else _reduce(loss, reduction)


def _l2(inp, target, reduction):
    """L2Loss::forward (tch-modules/src/l2_loss.rs:13-15): mse_loss on the logits."""
    return F.mse_loss(inp, target, reduction=reduction)


def _cross_entropy(inp, target_dense, reduction):
    """CrossEntropyLoss::new(false, ..)::forward (tch-modules/src/cross_entropy.rs:19-58): arg-max of
    the dense target, then cross_entropy_for_logits = log_softmax + nll per row."""
    sparse = target_dense.max(dim=1)[1]
    if inp.numel() == 0 and reduction == "mean":
        return torch.zeros((), dtype=torch.float32, device=inp.device)
    # tch 0.7 `cross_entropy_for_logits` = log_softmax(-1).nll_loss(target) and `nll_loss` is the
    # generated wrapper with Reduction::Mean: ALREADY a scalar mean over the rows, so the
    # .sum() / .mean() of cross_entropy.rs:47-51 change nothing (M = 0 with Sum: NaN, like ATen)
    loss = F.nll_loss(F.log_softmax(inp, dim=-1), sparse, reduction="mean")
    return loss if reduction == "none" else _reduce(loss, reduction)


def _kind_loss(kind, inp, target, pos_weight, gamma, reduction):
    kind = kind.lower()
    if kind == "bce":
        return _bce(inp, target, pos_weight, reduction)
    if kind == "focal":
        return _focal(inp, target, pos_weight, gamma, reduction)
    if kind == "l2":
        return _l2(inp, target, reduction)
    if kind == "crossentropy":
        return _cross_entropy(inp, target, reduction)
    raise ValueError(kind)


def yolo_loss_from_matches(merged: dict, batches, flats, target_cycxhw, target_class,
                           cfg: LossConfig):
    """loss_.rs:236-273 given the matcher's output.  Returns (losses dict, iou_score, pred rows)."""
    dev = merged["cy"].device
    B = merged["cy"].shape[0]
    pred = odet.index_by_flats(merged, batches, flats)  # pred_target_matching.rs:242
    M = batches.shape[0]
    p = (pred["cy"], pred["cx"], pred["h"], pred["w"])
    q = tuple(target_cycxhw[:, i:i + 1] for i in range(4))

    # iou_loss (:279-322)
    if cfg.box_metric == "hausdorff":  # (:285-288): the distance is the loss, no score
        iou_score = None
        iou_loss = obox.hausdorff_distance_to(p, q)
    else:
        iou_score = obox.METRICS[cfg.box_metric](p, q)
        iou_loss = 1.0 - iou_score
    per_row = iou_loss
    if cfg.reduction == "mean":
        iou_loss = iou_loss.mean() if M > 0 else torch.zeros((), dtype=torch.float32, device=dev)
    else:
        iou_loss = iou_loss.sum()

    # classification_loss (:324-374)
    pos = 1.0 - 0.5 * cfg.smooth_classification_coef
    neg = 1.0 - pos
    pred_class = pred["class_logit"]
    with torch.no_grad():
        sparse = target_class.view(-1, 1)
        pos_values = torch.full(sparse.shape, pos, dtype=torch.float32, device=dev)
        target_dense = torch.full_like(pred_class, neg).scatter_(1, sparse, pos_values)
    cls_loss = _kind_loss(cfg.classification_loss_kind, pred_class, target_dense, None,
                          cfg.focal_loss_gamma, cfg.reduction)

    # objectness_loss (:376-468)
    pred_obj = merged["obj_logit"]
    with torch.no_grad():
        target_scores = torch.full((M, 1), 1.0 - cfg.smooth_objectness_coef, dtype=torch.float32,
                                   device=dev)
        if iou_score is not None:  # (:401-404)
            target_scores += iou_score.detach().clamp(0.0, 1.0) * cfg.smooth_objectness_coef
        target_obj = torch.zeros_like(pred_obj)
        # index_put_(&[Some(batches), None, Some(flats)], target_scores, false) (:409-414)
        target_obj[batches, :, flats] = target_scores
    pw = None
    if cfg.objectness_pos_weight is not None:
        pw = torch.tensor([cfg.objectness_pos_weight], dtype=torch.float32, device=dev)
    obj_loss = _kind_loss(cfg.objectness_loss_kind, pred_obj.view(B, -1), target_obj.view(B, -1), pw,
                          cfg.focal_loss_gamma, cfg.reduction)

    total = (cfg.iou_loss_weight * iou_loss + cfg.classification_loss_weight * cls_loss
             + cfg.objectness_loss_weight * obj_loss)  # (:258-260)
    losses = dict(total=total, iou=iou_loss, classification=cls_loss, objectness=obj_loss)
    if iou_score is None:  # callers compare per-row values: hand back the distances
        iou_score = per_row
    return losses, iou_score, pred, target_dense, target_obj


def yolo_loss(heads: Sequence[torch.Tensor], infos, num_classes: int, targets, cfg: LossConfig,
              want_grad: bool = True, use_c_matcher: bool = False):
    """Detect2D x L -> merge -> match -> loss (-> autograd backward to the head tensors),
    i.e. what single_gpu.rs:152-168 runs after the backbone."""
    heads = [h.detach().clone().requires_grad_(want_grad) for h in heads]
    merged = odet.forward_merged(heads, infos, num_classes)
    mfn = omatch.match_bboxes_c if use_c_matcher else omatch.match_bboxes
    mb, mf, mg, mt, mc = mfn(infos, targets, cfg.match_rect4, cfg.anchor_scale_thresh)
    dev = heads[0].device
    batches = torch.from_numpy(mb).to(dev)
    flats = torch.from_numpy(mf).to(dev)
    tgt = torch.from_numpy(mt).to(dev)
    cls = torch.from_numpy(mc).to(dev)
    losses, iou_score, pred, _, _ = yolo_loss_from_matches(merged, batches, flats, tgt, cls, cfg)
    grads = None
    if want_grad:
        if losses["total"].requires_grad:
            losses["total"].backward()
        grads = [h.grad if h.grad is not None else torch.zeros_like(h) for h in heads]
    return dict(losses={k: v.detach() for k, v in losses.items()}, iou_score=iou_score.detach(),
                pred={k: v.detach() for k, v in pred.items()}, grads=grads,
                matches=(mb, mf, mg, mt, mc))
