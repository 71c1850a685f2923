"""NonMaxSuppression::forward / YoloInference::forward restated
(yolo-dl/src/loss/nms.rs:72-170, yolo-dl/src/loss/inference.rs:37-93).

`nms_by_scores` stands in for the un-vendored `tch_nms::nms_by_scores` (tch-nms @138b5e40,
yolo-dl/Cargo.toml:48; call site nms.rs:142-148).  Its published contract is "grouped NMS by
scores"; the algorithm restated here is the standard greedy one that torchvision's
`batched_nms` implements (cross-checked in tests/test_oracle.py): candidates in descending
score order, equal scores in ascending row order, a candidate is dropped iff an already kept
box of the same group has IoU > threshold, IoU = inter / (area_a + area_b - inter) in f32.
PARITY UNPINNED at this boundary (no reference test or source).

Test infrastructure only (see oracle/__init__.py).
"""
import numpy as np
import torch


def nms_by_scores(ltrb: np.ndarray, scores: np.ndarray, groups: np.ndarray, iou_threshold: float):
    """Greedy grouped NMS.  ltrb [K,4] f32 (l,t,r,b).  Returns kept row indices, score-descending
    (ties: lower row first)."""
    ltrb = np.asarray(ltrb, dtype=np.float32).reshape(-1, 4)
    scores = np.asarray(scores, dtype=np.float32).reshape(-1)
    groups = np.asarray(groups, dtype=np.int64).reshape(-1)
    K = scores.shape[0]
    if K == 0:
        return np.zeros(0, dtype=np.int64)
    order = np.lexsort((np.arange(K), -scores.astype(np.float64)))  # stable: score desc, row asc
    x1, y1, x2, y2 = (ltrb[order, i] for i in range(4))
    grp = groups[order]
    areas = (x2 - x1) * (y2 - y1)  # f32
    thr = float(iou_threshold)
    # incremental: test candidate i against the kept set (vectorised over kept)
    kept_pos = np.zeros(K, dtype=np.int64)
    nk = 0
    for i in range(K):
        if nk:
            kp = kept_pos[:nk]
            same = grp[kp] == grp[i]
            if same.any():
                kp = kp[same]
                xx1 = np.maximum(x1[kp], x1[i])
                yy1 = np.maximum(y1[kp], y1[i])
                xx2 = np.minimum(x2[kp], x2[i])
                yy2 = np.minimum(y2[kp], y2[i])
                w = np.maximum(np.float32(0), xx2 - xx1)
                h = np.maximum(np.float32(0), yy2 - yy1)
                inter = w * h
                with np.errstate(divide="ignore", invalid="ignore"):
                    ovr = inter / (areas[kp] + areas[i] - inter)
                if (ovr.astype(np.float64) > thr).any():
                    continue
        kept_pos[nk] = i
        nk += 1
    return order[kept_pos[:nk]].astype(np.int64)


def nms_forward(merged: dict, conf_thr: float, iou_thr: float, by_class: bool,
                compat_chunk65536: bool = False, use_torchvision: bool = False):
    """NonMaxSuppression::forward (nms.rs:72-170) on a merged detection (dict of torch tensors).

    Returns dict(batches, classes, instances i64 [K]; tlbr f32 [K,4]; conf f32 [K]) with kept
    rows ordered by (batch, score desc, nonzero-order asc) — the order of yh_nms.
    The default is one global NMS; `compat_chunk65536` reproduces the reference's independent
    65536-row chunks (:135-151) but WITH the chunk offset added (the reference omits it, :151).
    """
    with torch.no_grad():
        C = merged["class_logit"].shape[1]
        cy, cx, h, w = merged["cy"], merged["cx"], merged["h"], merged["w"]
        obj_prob = merged["obj_logit"].sigmoid()
        confidence = obj_prob * merged["class_logit"].sigmoid()  # merged_dense_detection.rs:143-153
        t = cy - h / 2.0
        b = cy + h / 2.0
        l = cx - w / 2.0
        r = cx + w / 2.0
        mask = obj_prob.ge(conf_thr).logical_and(confidence.ge(conf_thr))
        idx = mask.nonzero()
        batches, classes, instances = idx[:, 0], idx[:, 1], idx[:, 2]
        new_t = t[batches, :, instances]
        new_l = l[batches, :, instances]
        new_b = b[batches, :, instances]
        new_r = r[batches, :, instances]
        conf = confidence[batches, classes, instances].view(-1, 1)
        K = batches.shape[0]
        ltrb = torch.cat([new_l, new_t, new_r, new_b], 1)
        group = batches * C + classes if by_class else batches.clone()
        # everything up to here is the reference's ATen op sequence and runs on whatever device
        # `merged` lives on (CPU, or CUDA = the reference's real deployment through libtorch);
        # the greedy stand-in for tch-nms works on host copies of its inputs, torchvision's
        # batched_nms (use_torchvision) on the tensors where they are
        chunk = 65536 if compat_chunk65536 else max(K, 1)
        tv_keeps = None
        if use_torchvision and K > 0:
            import torchvision
            tv_keeps = [torchvision.ops.batched_nms(ltrb[s:s + chunk], conf[s:s + chunk].view(-1),
                                                    group[s:s + chunk], float(iou_thr)).cpu().numpy() + s
                        for s in range(0, K, chunk)]
        tlbr_t = torch.cat([new_t, new_l, new_b, new_r], 1).cpu()
        ltrb, conf, group = ltrb.cpu(), conf.cpu(), group.cpu()
        batches, classes, instances = batches.cpu(), classes.cpu(), instances.cpu()
        if K == 0:
            keep = np.zeros(0, dtype=np.int64)
        elif tv_keeps is not None:
            keep = np.concatenate(tv_keeps)
        else:
            keep = np.concatenate([nms_by_scores(ltrb[s:s + chunk].numpy(), conf[s:s + chunk].view(-1).numpy(),
                                                 group[s:s + chunk].numpy(), iou_thr) + s
                                   for s in range(0, K, chunk)])
        kb = batches.numpy()[keep]
        kc = classes.numpy()[keep]
        ki = instances.numpy()[keep]
        kconf = conf.view(-1).numpy()[keep]
        N = merged["cy"].shape[2]
        order = np.lexsort((kc * N + ki, -kconf.astype(np.float64), kb))
        keep = keep[order]
        tlbr = tlbr_t.numpy()
        return dict(batches=batches.numpy()[keep], classes=classes.numpy()[keep],
                    instances=instances.numpy()[keep], tlbr=tlbr[keep],
                    conf=conf.view(-1).numpy()[keep], num_candidates=K,
                    cand=(batches.numpy(), classes.numpy(), instances.numpy()))


def inference_pick(out: dict):
    """YoloInference::forward post-pass (inference.rs:50-77): per (batch, instance) keep the row
    with max confidence; Iterator::max_by_key returns the LAST maximum."""
    best = {}
    for i, (b, n, c) in enumerate(zip(out["batches"], out["instances"], out["conf"])):
        k = (int(b), int(n))
        if k not in best or not (c < out["conf"][best[k]]):  # >= : later row wins ties
            best[k] = i
    sel = np.array(sorted(best.values()), dtype=np.int64)
    return {k: (v[sel] if isinstance(v, np.ndarray) else v) for k, v in out.items()}
