"""CyCxHWMatcher::match_bboxes restated (yolo-dl/src/loss/pred_target_matching.rs:51-276).

Host f64 logic; Python floats are IEEE doubles and Python never fuses a*b+c, like the Rust.
`match_bboxes` is the readable restatement (insertion-ordered dict instead of the reference's
HashMap, so the order is deterministic); `match_bboxes_c` runs the same fold in C
(oracle/matcher.c, built by oracle/build.py) and is what the CPU baseline times.

Test infrastructure only (see oracle/__init__.py).
"""
import ctypes
import math
import os
from typing import List, Sequence

import numpy as np

F64_EPS = 2.220446049250313e-16  # approx::abs_diff_eq! default epsilon for f64 (:65)


def match_bboxes(infos, targets: Sequence[Sequence[Sequence[float]]], rect4: bool = True,
                 anchor_scale_thresh: float = 4.0):
    """targets[b] = list of (cy, cx, h, w, class) in ratio units.

    Returns (batches, flats, gt_index_in_image, target_rows) sorted by (batch, flat); the
    reference's own order is HashMap order, so sets are what is comparable.
    """
    assert anchor_scale_thresh >= 1.0, "anchor_scale_thresh must be greater than or equal to 1"
    snap = 0.5
    matchings = {}
    for b, boxes in enumerate(targets):
        for gi, (gcy, gcx, gh, gw, gcls) in enumerate(boxes):
            if abs(gh - 0.0) <= F64_EPS or abs(gw - 0.0) <= F64_EPS:  # (:62-73)
                continue
            for li, info in enumerate(infos):
                H, W = info.H, info.W
                # Transform::from_sizes_exact(HW::unit(), feature) (bbox/src/transform.rs:16-37):
                # src tlbr = (0,0,0+1,0+1), tgt = (0,0,0+H,0+W); sy = tgt.h/src.h; ty = tgt.t - src.t*sy
                sy = ((0.0 + float(H)) - 0.0) / ((0.0 + 1.0) - 0.0)
                sx = ((0.0 + float(W)) - 0.0) / ((0.0 + 1.0) - 0.0)
                ty = 0.0 - 0.0 * sy
                tx = 0.0 - 0.0 * sx
                cy = gcy * sy + ty  # CyCxHW::transform (bbox/src/cycxhw.rs:40-47)
                cx = gcx * sx + tx
                row = int(math.floor(cy))
                col = int(math.floor(cx))
                cy_fract = cy - math.trunc(cy)
                cx_fract = cx - math.trunc(cx)
                cells = [(row, col)]
                if cy_fract < snap:
                    cells.append((row - 1, col))
                if cx_fract < snap:
                    cells.append((row, col - 1))
                if rect4 and cy_fract > 1.0 - snap:
                    cells.append((row + 1, col))
                if rect4 and cx_fract > 1.0 - snap:
                    cells.append((row, col + 1))
                cells = [(r, c) for r, c in cells if 0 <= r < H and 0 <= c < W]  # (:112-116)
                for ai, (ah, aw) in enumerate(info.anchors):
                    bounded = (gh / ah <= anchor_scale_thresh and ah / gh <= anchor_scale_thresh
                               and gw / aw <= anchor_scale_thresh and aw / gw <= anchor_scale_thresh)
                    if not bounded:  # (:139-150)
                        continue
                    for r, c in cells:
                        key = (b, li, ai, r, c)
                        if key in matchings:  # (:185-210)
                            ogi = matchings[key]
                            ocy, ocx = targets[b][ogi][0], targets[b][ogi][1]
                            pcy = (float(r) + 0.5) / float(H)
                            pcx = (float(c) + 0.5) / float(W)
                            d_orig = (ocy - pcy) * (ocy - pcy) + (ocx - pcx) * (ocx - pcx)
                            d_new = (gcy - pcy) * (gcy - pcy) + (gcx - pcx) * (gcx - pcx)
                            if d_new < d_orig:
                                matchings[key] = gi
                        else:
                            matchings[key] = gi
    rows = []
    for (b, li, ai, r, c), gi in matchings.items():
        info = infos[li]
        flat = info.flat_start + c + info.W * (r + info.H * ai)  # merged_dense_detection.rs:374-376
        rows.append((b, flat, gi))
    rows.sort()
    batches = np.array([x[0] for x in rows], dtype=np.int64)
    flats = np.array([x[1] for x in rows], dtype=np.int64)
    gts = np.array([x[2] for x in rows], dtype=np.int64)
    tgt = np.array([targets[b][g][:4] for b, _, g in rows], dtype=np.float64).reshape(-1, 4)
    cls = np.array([int(targets[b][g][4]) for b, _, g in rows], dtype=np.int64)
    # label.rect.cast::<f32>() (:247)
    return batches, flats, gts, tgt.astype(np.float32), cls


# ----------------------------------------------------------------------------------
# C version (same fold, open-addressing hash map)
# ----------------------------------------------------------------------------------
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "liboracle.so")
        if not os.path.exists(path):
            from . import build
            build.build()
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_match_bboxes.restype = ctypes.c_long
    return _LIB


def pack_targets(targets):
    """-> (gt [G,5] f64, offsets [B+1] i32)."""
    offs = np.zeros(len(targets) + 1, dtype=np.int32)
    for i, t in enumerate(targets):
        offs[i + 1] = offs[i] + len(t)
    gt = np.array([row for t in targets for row in t], dtype=np.float64).reshape(-1, 5)
    return np.ascontiguousarray(gt), offs


def match_bboxes_c(infos, targets, rect4=True, anchor_scale_thresh=4.0):
    gt, offs = pack_targets(targets)
    L = len(infos)
    Hs = np.array([i.H for i in infos], dtype=np.int32)
    Ws = np.array([i.W for i in infos], dtype=np.int32)
    As = np.array([i.A for i in infos], dtype=np.int32)
    starts = np.array([i.flat_start for i in infos], dtype=np.int64)
    amax = max(i.A for i in infos)
    anch = np.zeros((L, amax, 2), dtype=np.float64)
    for l, i in enumerate(infos):
        anch[l, : i.A] = np.array(i.anchors, dtype=np.float64)
    cap = max(1, gt.shape[0] * sum(i.A for i in infos) * 5)
    ob = np.empty(cap, dtype=np.int64)
    of = np.empty(cap, dtype=np.int64)
    og = np.empty(cap, dtype=np.int64)
    P = ctypes.c_void_p
    m = _lib().oracle_match_bboxes(
        P(gt.ctypes.data), P(offs.ctypes.data), ctypes.c_int(len(targets)), ctypes.c_int(L),
        P(Hs.ctypes.data), P(Ws.ctypes.data), P(As.ctypes.data), P(starts.ctypes.data),
        P(anch.ctypes.data), ctypes.c_int(amax), ctypes.c_int(1 if rect4 else 0),
        ctypes.c_double(anchor_scale_thresh), P(ob.ctypes.data), P(of.ctypes.data),
        P(og.ctypes.data), ctypes.c_long(cap))
    assert 0 <= m <= cap
    ob, of, og = ob[:m], of[:m], og[:m]
    order = np.lexsort((of, ob))
    ob, of, og = ob[order], of[order], og[order]
    rows = gt[offs[ob] + og] if m else np.zeros((0, 5))
    return ob, of, og, rows[:, :4].astype(np.float32), rows[:, 4].astype(np.int64)
