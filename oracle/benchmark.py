"""YoloBenchmark::forward restated with the reference's ATen op order
(yolo-dl/src/loss/benchmark.rs:34-100).  Test infrastructure only (see oracle/__init__.py)."""
import torch


def yolo_benchmark(merged: dict, pred: dict, target_class: torch.Tensor, confidence_threshold: float):
    """merged: dict of the merged detection (obj_logit [B,1,N] ...); pred: gathered matched rows
    (obj_logit [M,1], class_logit [M,C]); target_class [M] or [M,1] i64.
    Returns (dict of the four f64 ratios, the five integer counts)."""
    thr = confidence_threshold
    with torch.no_grad():
        # objectness (:43-79)
        all_mask = merged["obj_logit"].sigmoid().ge(thr)
        matched_obj = pred["obj_logit"].sigmoid()
        matched_mask = matched_obj.ge(thr)
        all_count = all_mask.numel()
        all_pos = int(all_mask.count_nonzero())
        all_neg = all_count - all_pos
        matched_count = matched_mask.numel()
        matched_pos = int(matched_mask.count_nonzero())
        matched_neg = matched_count - matched_pos
        unmatched_neg = all_neg - matched_neg
        accuracy = (matched_pos + unmatched_neg) / all_count
        recall = matched_pos / matched_count if matched_count != 0 else 1.0
        precision = matched_pos / all_pos if all_pos != 0 else 1.0
        # classification (:82-94)
        class_hits = 0
        if pred["class_logit"].numel() != 0:
            conf = matched_obj * pred["class_logit"].sigmoid()          # ObjectDetectionTensor::confidence
            conf_mask = conf.ge(thr)
            _, pred_class = pred["class_logit"].max(1, keepdim=True)
            class_mask = target_class.view(-1, 1).eq(pred_class)
            mask = conf_mask.any(1, keepdim=True).logical_and(class_mask)
            class_hits = int(mask.count_nonzero())
            class_accuracy = class_hits / mask.numel()
        else:
            class_accuracy = 1.0
    return (dict(obj_accuracy=accuracy, obj_recall=recall, obj_precision=precision,
                 class_accuracy=class_accuracy),
            [all_count, all_pos, matched_count, matched_pos, class_hits])
