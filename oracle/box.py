"""Batched box metrics of tch-goodies::compound_tensor, same ATen op order.

Test infrastructure only (see oracle/__init__.py).  Boxes are dicts/tuples of [M,1] tensors.
"""
import math

import torch

EPSILON = 1e-16  # tch-goodies/src/utils.rs:5


def tlbr_of(cy, cx, h, w):
    """From<&CyCxHWTensor> for TLBRTensor (tlbr.rs:214-225)."""
    t = cy - h / 2.0
    b = cy + h / 2.0
    l = cx - w / 2.0
    r = cx + w / 2.0
    return t, l, b, r


def intersect_area(a, b):
    """TLBRTensor::intersect_area_with (tlbr.rs:81-108)."""
    at, al, ab, ar = a
    bt, bl, bb, br = b
    max_t = torch.maximum(at, bt)
    max_l = torch.maximum(al, bl)
    min_b = torch.minimum(ab, bb)
    min_r = torch.minimum(ar, br)
    inner_h = (min_b - max_t).clamp_min(0.0)
    inner_w = (min_r - max_l).clamp_min(0.0)
    return inner_h * inner_w


def closure(a, b):
    """TLBRTensor::closure_with (tlbr.rs:111-134)."""
    at, al, ab, ar = a
    bt, bl, bb, br = b
    return torch.minimum(at, bt), torch.minimum(al, bl), torch.maximum(ab, bb), torch.maximum(ar, br)


def iou_with(p, q):
    """CyCxHWTensor::iou_with (cycxhw.rs:67-72). p, q = (cy,cx,h,w)."""
    inter = intersect_area(tlbr_of(*p), tlbr_of(*q))
    outer = p[2] * p[3] + q[2] * q[3] - inter + EPSILON
    return inter / outer


def giou_with(p, q):
    """cycxhw.rs:75-82."""
    inter = intersect_area(tlbr_of(*p), tlbr_of(*q))
    outer = p[2] * p[3] + q[2] * q[3] - inter + EPSILON
    ct, cl, cb, cr = closure(tlbr_of(*p), tlbr_of(*q))
    # closure converted back to CyCxHW (h=b-t, w=r-l), area = h*w
    closure_area = (cb - ct) * (cr - cl)
    iou = inter / outer
    return iou - (closure_area - outer) / (closure_area + EPSILON)


def _diag_and_dist(p, q):
    ct, cl, cb, cr = closure(tlbr_of(*p), tlbr_of(*q))
    ch, cw = cb - ct, cr - cl
    diagonal_square = ch.pow(2) + cw.pow(2) + EPSILON
    center_dist_square = (p[0] - q[0]).pow(2) + (p[1] - q[1]).pow(2)
    return diagonal_square, center_dist_square


def diou_with(p, q):
    """cycxhw.rs:85-97."""
    iou = iou_with(p, q)
    diagonal_square, center_dist_square = _diag_and_dist(p, q)
    return iou - center_dist_square / diagonal_square


def ciou_with(p, q):
    """cycxhw.rs:100-120 — note the '+' in front of alpha*v (:119)."""
    iou = iou_with(p, q)
    pred_angle = torch.atan2(p[2], p[3])
    target_angle = torch.atan2(q[2], q[3])
    diagonal_square, center_dist_square = _diag_and_dist(p, q)
    shape_loss = (pred_angle - target_angle).pow(2) * 4.0 / math.pi ** 2
    with torch.no_grad():
        shape_loss_coef = shape_loss / (1.0 - iou + shape_loss)
    return iou - center_dist_square / diagonal_square + shape_loss_coef * shape_loss


def hausdorff_distance_to(p, q):
    """CyCxHWTensor::hausdorff_distance_to (cycxhw.rs:122-125) -> TLBRTensor::hausdorff_distance_to
    (tlbr.rs:137-177): self = p, other = q; `pow(&Tensor::from(2i64))` = pow with a 0-dim int64
    tensor exponent; `max_other` = torch.maximum."""
    tl, ll, bl, rl = tlbr_of(*p)
    tr, lr, br, rr = tlbr_of(*q)
    dt = tr - tl
    dl = lr - ll
    db = bl - br
    dr = rl - rr
    dt_l, dl_l, db_l, dr_l = (x.clamp_min(0.0) for x in (dt, dl, db, dr))
    dt_r, dl_r, db_r, dr_r = ((-x).clamp_min(0.0) for x in (dt, dl, db, dr))
    two = torch.tensor(2, dtype=torch.int64)
    out = dt_l.pow(two) + dl_l.pow(two)
    for a, b in ((dt_l, dr_l), (db_l, dl_l), (db_l, dr_l), (dt_r, dl_r), (dt_r, dr_r), (db_r, dl_r),
                 (db_r, dr_r)):
        out = torch.maximum(out, a.pow(two) + b.pow(two))
    return out.sqrt()


METRICS = {"iou": iou_with, "giou": giou_with, "diou": diou_with, "ciou": ciou_with}
