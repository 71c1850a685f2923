"""GPU parity: NonMaxSuppression / YoloInference / nms_by_scores (yh_nms, yh_inference_pick,
yh_nms_by_scores) against the oracle.

Index parity is bit-exact.  The candidate set depends on `sigmoid(obj) >= thr` and
`conf >= thr`, where the device's expf and the CPU's vectorised exp may differ in the last
bit; `_stable` asserts that no candidate of the seeded input sits within 4 ulp of the
threshold, so that the exact comparison is meaningful (it never triggers for these seeds).
"""
import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh
from oracle import detect as odet
from oracle import nms as onms
from yolo_dl_b200 import synth

from util import build, close, rows_set

pytestmark = pytest.mark.gpu


def _stable(merged_ref, thr):
    obj = merged_ref["obj_logit"].sigmoid()
    conf = obj * merged_ref["class_logit"].sigmoid()
    eps = 4 * 6e-8 * max(thr, 1e-3)
    near = ((obj - thr).abs() < eps).sum() + (((conf - thr).abs() < eps) & (obj >= thr)).sum()
    assert int(near) == 0, "seeded input has a borderline candidate; pick another seed"


def _check_nms(dev, spec, B, seed, conf_thr, iou_thr, by_class, obj_mean=-4.0, cls_mean=-3.0,
               cand=8192, mutate=None, repeat=1):
    heads_cpu = synth.make_heads(spec, B, seed, obj_mean=obj_mean, cls_mean=cls_mean)
    if mutate is not None:
        mutate(heads_cpu)
    heads, _, merged, infos = build(spec, heads_cpu, dev)
    nms = yh.NonMaxSuppressionInit(iou_thr, conf_thr, by_class, max_candidates_per_image=cand).build()
    for _ in range(repeat):  # the same module (workspace) again: the last result is compared
        out = nms.forward(merged)
    mref = odet.forward_merged(heads_cpu, infos, spec.num_classes)
    _stable(mref, conf_thr)
    ref = onms.nms_forward(mref, conf_thr, iou_thr, by_class)
    got = (out.batches.cpu().numpy(), out.classes.cpu().numpy(), out.instances.cpu().numpy())
    # bit-exact kept rows, in the documented order (batch, score desc, nonzero index asc)
    assert rows_set(*got) == rows_set(ref["batches"], ref["classes"], ref["instances"])
    conf_equal = np.array_equal(out.confidence.view(-1).cpu().numpy(), ref["conf"])
    if conf_equal:  # identical scores => identical order
        assert np.array_equal(got[0], ref["batches"]) and np.array_equal(got[2], ref["instances"])
        assert np.array_equal(got[1], ref["classes"])
    # values within 1e-5 relative, matched row by row through the (b,c,n) key
    N = merged.num_instances()
    key_g = (got[0] * spec.num_classes + got[1]) * N + got[2]
    key_r = (ref["batches"] * spec.num_classes + ref["classes"]) * N + ref["instances"]
    og, orr = np.argsort(key_g), np.argsort(key_r)
    b = out.bbox
    tl = torch.cat([b.t, b.l, b.b, b.r], 1).cpu().numpy()
    close(tl[og], ref["tlbr"][orr], what="tlbr")
    close(out.confidence.view(-1).cpu().numpy()[og], ref["conf"][orr], what="conf")
    return out, ref


@pytest.mark.parametrize("by_class", [False, True])
@pytest.mark.parametrize("conf_thr,iou_thr", [(0.4, 0.5), (0.1, 0.6)])
def test_nms_yolov4_608_batch8(dev, by_class, conf_thr, iou_thr):
    """BASELINE config 0: YOLOv4-608 decode+NMS, batch 8, shipped (.4/.5) and default (.1/.6)
    thresholds, per image and per class."""
    out, ref = _check_nms(dev, synth.yolov4_608(), 8, synth.SEED0, conf_thr, iou_thr, by_class)
    assert out.num_samples() > 0


@pytest.mark.parametrize("by_class", [False, True])
def test_nms_tiny416(dev, by_class):
    """BASELINE config 2 (batch reduced to 32 for the CPU oracle; 256 is covered by properties)."""
    _check_nms(dev, synth.yolov4_tiny_416(), 32, synth.SEED0 + 2, 0.1, 0.6, by_class)


@pytest.mark.parametrize("by_class", [False, True])
def test_nms_small_dense(dev, by_class):
    """a head where most cells pass: many candidates per image, several sort tiles."""
    _check_nms(dev, synth.small_spec(), 3, 5, 0.25, 0.45, by_class, obj_mean=1.0, cls_mean=0.5)


@pytest.mark.parametrize("by_class", [False, True])
@pytest.mark.parametrize("name,B", [("yolov4-608", 8), ("yolov4-tiny-416", 32), ("yolov4-tiny-416", 160)])
def test_nms_small_capacity_emit_paths(dev, name, B, by_class):
    """Candidate capacity <= 1024 (what a training step uses): no large-segment kernel, so
    nms_segment_kernel emits the kept lists itself while every image's CTA is resident (B <= SM
    count: look-back over the kept counts its predecessors publish) and the separate emit kernel
    runs above that (160 images).  Two images in the middle of the batch have no candidate at all
    (their CTAs publish a count of zero); called twice: the second call starts from the words the
    first one left."""
    spec = synth.SPECS[name]()

    def mute(heads):  # objectness far below any threshold for two images
        for h in heads:
            E = 5 + spec.num_classes
            v = h.view(h.shape[0], E, -1)
            v[B // 2, 4] = -30.0
            v[B // 2 + 1, 4] = -30.0

    out, ref = _check_nms(dev, spec, B, 20261108, 0.4, 0.5, by_class, cand=1024, mutate=mute, repeat=2)
    got_b = out.batches.cpu().numpy()
    assert out.num_samples() > 0 and not np.isin(got_b, [B // 2, B // 2 + 1]).any()


def test_nms_no_candidates(dev):
    out, ref = _check_nms(dev, synth.small_spec(), 2, 5, 0.999, 0.5, False, obj_mean=-8.0)
    assert out.num_samples() == 0 and ref["num_candidates"] == 0


def test_nms_candidate_overflow_is_reported(dev):
    spec = synth.small_spec()
    heads_cpu = synth.make_heads(spec, 2, 5, obj_mean=3.0, cls_mean=3.0)
    _, _, merged, _ = build(spec, heads_cpu, dev)
    nms = yh.NonMaxSuppressionInit(0.5, 0.1, False, max_candidates_per_image=64, grow=False).build()
    with pytest.raises(yh.YoloHotError, match="overflow") as e:
        nms.forward(merged)
    assert e.value.status == yh._lib.YH_ERR_CAPACITY
    # the filter counted past the capacity: the host learns how much room the batch needs
    need = yh.load().yh_workspace_candidates_needed(nms._ws.handle)
    assert need > 64
    # ... and with `grow` (the default) the same call succeeds: count -> allocate -> fill
    grown = yh.NonMaxSuppressionInit(0.5, 0.1, False, max_candidates_per_image=64).build()
    out = grown.forward(merged)
    assert grown.max_candidates_per_image >= need
    big = yh.NonMaxSuppressionInit(0.5, 0.1, False, max_candidates_per_image=1 << 16).build().forward(merged)
    assert torch.equal(out.instances, big.instances) and torch.equal(out.batches, big.batches)


@pytest.mark.parametrize("K", [0, 1, 2, 33, 1000, 5000, 8192, 8193, 20000, 65536, 70000])
@pytest.mark.parametrize("groups", [1, 80])
def test_nms_by_scores_exact(dev, K, groups):
    """drop-in for tch_nms::nms_by_scores: kept indices AND their order are bit-exact (no
    transcendental on this path: +,-,*,/ and min/max are IEEE-exact on both sides).
    Covers the shared-memory sort (<= 8192), the hybrid sort and > 65536 rows, 5 % duplicate
    scores (tie-break by index)."""
    if K > 20000 and groups == 80:
        pytest.skip("one group count is enough for the large sizes")
    ltrb, scores, grp = synth.make_nms_boxes(K, 100 + K, num_groups=groups)
    ref = onms.nms_by_scores(ltrb, scores, grp, 0.5)
    got = yh.nms_by_scores(torch.from_numpy(ltrb).to(dev), torch.from_numpy(scores).to(dev),
                           torch.from_numpy(grp).to(dev), 0.5)
    assert np.array_equal(got.cpu().numpy(), ref)


def test_nms_by_scores_matches_torchvision(dev):
    import torchvision
    ltrb, scores, grp = synth.make_nms_boxes(3000, 9, num_groups=7, dup_frac=0.0)
    ref = torchvision.ops.batched_nms(torch.from_numpy(ltrb), torch.from_numpy(scores),
                                      torch.from_numpy(grp), 0.6).numpy()
    got = yh.nms_by_scores(torch.from_numpy(ltrb).to(dev), torch.from_numpy(scores).to(dev),
                           torch.from_numpy(grp).to(dev), 0.6)
    assert np.array_equal(got.cpu().numpy(), ref)


@pytest.mark.parametrize("K,groups", [(100000, 80), (200000, 80), (100000, 4), (65536, 1), (200000, 1)])
def test_nms_sweep_100k_200k(dev, K, groups):
    """BASELINE config 4, the sizes above one 65 536-row chunk: one global NMS over all rows,
    checked against torchvision's batched_nms (C++ on the host; the numpy oracle is O(K*kept))."""
    import torchvision
    ltrb, scores, grp = synth.make_nms_boxes(K, 7 + K + groups, num_groups=groups, dup_frac=0.0)
    ref = torchvision.ops.batched_nms(torch.from_numpy(ltrb), torch.from_numpy(scores),
                                      torch.from_numpy(grp), 0.5).numpy()
    got = yh.nms_by_scores(torch.from_numpy(ltrb).to(dev), torch.from_numpy(scores).to(dev),
                           torch.from_numpy(grp).to(dev), 0.5).cpu().numpy()
    # the kept SET is exact; among the ~K^2/2^24 pairs of equal f32 scores torchvision's final
    # sort is not stable, ours is (score descending, then row ascending)
    assert np.array_equal(np.sort(got), np.sort(ref))
    sc = scores[got]
    assert np.all(sc[:-1] >= sc[1:])
    ties = sc[:-1] == sc[1:]
    assert np.all(got[:-1][ties] < got[1:][ties])


def test_nms_by_scores_degenerate_boxes(dev):
    """zero-area and identical boxes, negative scores, NaN IoU (0/0 is not > thr)."""
    ltrb = np.array([[0, 0, 0, 0], [0, 0, 0, 0], [0.1, 0.1, 0.5, 0.5], [0.1, 0.1, 0.5, 0.5],
                     [0.1, 0.1, 0.5, 0.5]], dtype=np.float32)
    scores = np.array([0.5, 0.5, -1.0, 0.9, 0.9], dtype=np.float32)
    grp = np.array([0, 0, 0, 0, 1], dtype=np.int64)
    ref = onms.nms_by_scores(ltrb, scores, grp, 0.5)
    got = yh.nms_by_scores(torch.from_numpy(ltrb).to(dev), torch.from_numpy(scores).to(dev),
                           torch.from_numpy(grp).to(dev), 0.5)
    assert np.array_equal(got.cpu().numpy(), ref)
    assert ref.tolist() == [3, 4, 0, 1]


@pytest.mark.parametrize("by_class", [False, True])
def test_inference_pick(dev, by_class):
    """YoloInference::forward: one class per (batch, instance), the max-confidence one."""
    spec = synth.small_spec()
    heads_cpu = synth.make_heads(spec, 3, 8, obj_mean=0.5, cls_mean=0.0)
    heads, _, merged, infos = build(spec, heads_cpu, dev)
    inf = yh.YoloInferenceInit(0.5, 0.2, by_class).build()
    out = inf.forward(merged)
    mref = odet.forward_merged(heads_cpu, infos, spec.num_classes)
    _stable(mref, 0.2)
    ref = onms.inference_pick(onms.nms_forward(mref, 0.2, 0.5, by_class))
    got = rows_set(out.batches.cpu().numpy(), out.classes.cpu().numpy(), out.instances.cpu().numpy())
    assert got == rows_set(ref["batches"], ref["classes"], ref["instances"])
    pairs = list(zip(out.batches.cpu().tolist(), out.instances.cpu().tolist()))
    assert len(pairs) == len(set(pairs))
    if by_class:
        full = yh.NonMaxSuppressionInit(0.5, 0.2, True).build().forward(merged)
        assert full.num_samples() > out.num_samples()  # the pick removed something
    sel = out.batch_select(1)
    assert bool((sel.batches == 1).all())


def test_nms_full_size_properties(dev):
    """BASELINE sizes (YOLOv4-608 batch 64; tiny-416 batch 256): output sorted by (batch, score
    desc); no kept pair of one image overlaps above the threshold; NMS of the kept set keeps
    everything (idempotence); boxes equal the decode planes bit for bit."""
    for spec, B in ((synth.yolov4_608(), 64), (synth.yolov4_tiny_416(), 256)):
        heads = synth.make_heads(spec, B, 123, device=dev)
        detects = [yh.Detect2DInit(spec.num_classes, a).build() for a in spec.anchors]
        merged = yh.MergedDenseDetection.from_heads(heads, detects)
        for by_class in (False, True):
            out = yh.NonMaxSuppressionInit(0.6, 0.1, by_class).build().forward(merged)
            b = out.batches
            c = out.confidence.view(-1)
            assert bool((b[1:] >= b[:-1]).all())
            same = b[1:] == b[:-1]
            assert bool((c[1:][same] <= c[:-1][same]).all())
            assert bool((c >= 0.1).all())
            bx = out.bbox
            ltrb = torch.cat([bx.l, bx.t, bx.r, bx.b], 1).contiguous()
            grp = b * spec.num_classes + out.classes if by_class else b
            keep = yh.nms_by_scores(ltrb, c, grp, 0.6)
            assert keep.numel() == out.num_samples()
            # boxes agree with the decode kernel's planes bit for bit
            cy = merged.cy[b, 0, out.instances]
            hh = merged.h[b, 0, out.instances]
            assert torch.equal(bx.t.view(-1), cy - hh / 2.0)
