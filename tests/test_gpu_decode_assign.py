"""GPU parity: Detect2D / merge (yh_decode, yh_decode_merged) and target assignment (yh_assign)
against the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh
from oracle import detect as odet
from oracle import matcher as omatch
from yolo_dl_b200 import synth

from util import build, close

pytestmark = pytest.mark.gpu

CASES = [("small", 3), ("yolov4-tiny-416", 4), ("yolov4-608", 2), ("yolov4-csp-640", 1)]


@pytest.mark.parametrize("name,B", CASES)
def test_decode_layers_match_oracle(dev, name, B):
    """Detect2D::forward per layer (drop-in mode): cy,cx,h,w within 1e-5 relative; obj/class
    logits are views of the input (detect_2d.rs:136-139)."""
    spec = synth.SPECS[name]()
    heads_cpu = synth.make_heads(spec, B, synth.SEED0 + 1)
    for l, hc in enumerate(heads_cpu):
        det = yh.Detect2DInit(spec.num_classes, spec.anchors[l]).build()
        hd = hc.to(dev)
        out = det.forward(hd)
        ref = odet.detect2d_forward(hc, spec.num_classes, spec.anchors[l])
        for k in ("cy", "cx", "h", "w"):
            close(getattr(out, k), ref[k], what=f"{name} layer {l} {k}")
        assert out.obj_logit.untyped_storage().data_ptr() == hd.untyped_storage().data_ptr()
        assert out.class_logit.untyped_storage().data_ptr() == hd.untyped_storage().data_ptr()
        assert torch.equal(out.obj_logit.cpu(), ref["obj_logit"])
        assert torch.equal(out.class_logit.cpu(), ref["class_logit"])


@pytest.mark.parametrize("name,B", CASES)
def test_decode_merged_matches_oracle_and_layers(dev, name, B):
    """fused decode+merge == per-layer decode + MergedDenseDetection::from_detection_tensors,
    bit for bit; == oracle within 1e-5; merge∘split = id (merged_dense_detection.rs:105-116)."""
    spec = synth.SPECS[name]()
    heads_cpu = synth.make_heads(spec, B, synth.SEED0 + 2)
    heads, detects, merged, infos = build(spec, heads_cpu, dev)
    layers = [d.forward(h) for d, h in zip(detects, heads)]
    merged2 = yh.MergedDenseDetection.from_detection_tensors(yh.MergeDetect2D().forward(layers))
    assert torch.equal(merged.boxes, merged2.boxes)
    ref = odet.forward_merged(heads_cpu, infos, spec.num_classes)
    for k in ("cy", "cx", "h", "w"):
        close(getattr(merged, k), ref[k], what=f"{name} merged {k}")
    assert torch.equal(merged.obj_logit.cpu(), ref["obj_logit"])
    assert torch.equal(merged.class_logit.cpu(), ref["class_logit"])
    assert [i.flat_start for i in merged.info] == [i.flat_start for i in infos]


def test_decode_debug_assert_property(dev):
    """detect_2d.rs:74-119: cy == (sigmoid*2 - 0.5 + row)/H up to abs_diff (1e-6 here)."""
    spec = synth.yolov4_608()
    heads_cpu = synth.make_heads(spec, 1, 7)
    heads, _, merged, infos = build(spec, heads_cpu, dev)
    for hc, info in zip(heads_cpu, infos):
        v = hc.view(1, -1, info.A, info.H, info.W).double()
        rows = torch.arange(info.H).view(1, 1, info.H, 1).double()
        cols = torch.arange(info.W).view(1, 1, 1, info.W).double()
        cy = ((v[:, 0].sigmoid() * 2 - 0.5 + rows) / info.H).reshape(1, -1)
        cx = ((v[:, 1].sigmoid() * 2 - 0.5 + cols) / info.W).reshape(1, -1)
        sl = slice(info.flat_start, info.flat_start + info.size)
        assert (merged.cy[:, 0, sl].cpu().double() - cy).abs().max() < 1e-6
        assert (merged.cx[:, 0, sl].cpu().double() - cx).abs().max() < 1e-6


def test_decode_strided_output_and_odd_batch_stride(dev):
    """yh_decode into a merged buffer with a batch stride that breaks 16-byte alignment."""
    spec = synth.small_spec()
    heads_cpu = synth.make_heads(spec, 2, 11)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    N = sum(i.size for i in infos)
    stride = N + 3
    buf = torch.full((4, 2, stride), -7.0, device=dev)
    lib = yh.load()
    for hc, info in zip(heads_cpu, infos):
        hd = hc.to(dev)
        lay = yh._lib.make_layers([odet.LayerInfo(info.H, info.W, info.anchors, 0)])
        ptrs = [buf[k].data_ptr() + 4 * info.flat_start for k in range(4)]
        yh._lib.check(lib.yh_decode(hd.data_ptr(), 2, spec.num_classes, lay, *ptrs, stride,
                                    torch.cuda.current_stream().cuda_stream))
    ref = odet.forward_merged(heads_cpu, infos, spec.num_classes)
    for i, k in enumerate(("cy", "cx", "h", "w")):
        close(buf[i, :, :N], ref[k][:, 0], what=k)
    assert bool((buf[:, :, N:] == -7.0).all())


def _check_assign(dev, spec, B, targets, rect4=True, thresh=4.0, use_c=False):
    heads_cpu = synth.make_heads(spec, B, 5)
    _, _, merged, infos = build(spec, heads_cpu, dev)
    m = yh.CyCxHWMatcherInit("Rect4" if rect4 else "Rect2", thresh).build()
    out = m.match_bboxes(merged, targets)
    fn = omatch.match_bboxes_c if use_c else omatch.match_bboxes
    rb, rf, rg, rt, rc = fn(infos, targets, rect4, thresh)
    # bit-exact indices (sorted by (batch, flat) on both sides)
    assert np.array_equal(out.pred_indexes.batches.cpu().numpy(), rb)
    assert np.array_equal(out.pred_indexes.flats.cpu().numpy(), rf)
    offs = np.concatenate([[0], np.cumsum([len(t) for t in targets])])
    assert np.array_equal(out.gt_rows.cpu().numpy().astype(np.int64), offs[rb] + rg)
    tg = out.target
    got = torch.cat([tg.cycxhw.cy, tg.cycxhw.cx, tg.cycxhw.h, tg.cycxhw.w], 1).cpu().numpy()
    assert np.array_equal(got, rt)  # f64 -> f32 cast, exact
    assert np.array_equal(tg.class_.view(-1).cpu().numpy(), rc)
    # the pred rows are the prediction gathered at the matched indices
    p = out.pred
    assert torch.equal(p.cycxhw.cy.view(-1), merged.cy[out.pred_indexes.batches, 0, out.pred_indexes.flats])
    assert torch.equal(p.obj_logit.view(-1), merged.obj_logit[out.pred_indexes.batches, 0, out.pred_indexes.flats])
    return out


@pytest.mark.parametrize("rect4", [True, False])
def test_assign_small_adversarial(dev, rect4):
    spec = synth.small_spec()
    targets = synth.make_targets(spec, 4, 21, n_max=12, clustered=0.5, adversarial=True)
    _check_assign(dev, spec, 4, targets, rect4=rect4)


def test_assign_empty_and_ragged(dev):
    spec = synth.small_spec()
    targets = [[], [(0.5, 0.5, 0.2, 0.2, 1)], [], [(0.3, 0.7, 0.0, 0.1, 0)]]  # incl. zero-sized
    out = _check_assign(dev, spec, 4, targets)
    assert out.num_samples() > 0
    out = _check_assign(dev, spec, 2, [[], []])
    assert out.num_samples() == 0


@pytest.mark.parametrize("thresh", [1.0, 2.5, 4.0])
def test_assign_yolov4_608(dev, thresh):
    spec = synth.yolov4_608()
    targets = synth.make_targets(spec, 8, synth.SEED0 + 1)
    _check_assign(dev, spec, 8, targets, thresh=thresh, use_c=True)


def test_assign_csp640_dense_scene_300_gt(dev):
    """BASELINE config 3: 300 GT/image, 30 % clustered within 2 cells, adversarial centres."""
    spec = synth.yolov4_csp_640()
    targets = synth.make_targets(spec, 4, synth.SEED0 + 3, n_min=300, n_max=300, clustered=0.3,
                                 adversarial=True)
    out = _check_assign(dev, spec, 4, targets)
    # python and C matcher agree as well
    infos = odet.make_infos(spec.shapes, spec.anchors)
    a = omatch.match_bboxes(infos, targets)
    b = omatch.match_bboxes_c(infos, targets)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    # pred_target_matching.rs:219-235: matched cell within 1.5 cells of the GT centre
    bt, fl = out.pred_indexes.batches.cpu().numpy(), out.pred_indexes.flats.cpu().numpy()
    tcy = out.target.cycxhw.cy.view(-1).cpu().numpy().astype(np.float64)
    for info in infos:
        sel = (fl >= info.flat_start) & (fl < info.flat_start + info.size)
        rem = fl[sel] - info.flat_start
        row = (rem % (info.H * info.W)) // info.W
        assert np.all(np.abs(tcy[sel] * info.H - row) <= 1.5 + 1e-4)
    assert bt.max() == 3


def test_assign_rejects_bad_threshold(dev):
    with pytest.raises(ValueError, match="anchor_scale_thresh must be greater than or equal to 1"):
        yh.CyCxHWMatcherInit("Rect4", 0.5).build()


def test_decode_long_and_thin_maps(dev):
    """maps longer than the kernel's 1024-entry offset tables (per-element division instead) and
    planes whose size is not a multiple of 4 / of the 1024-cell chunk"""
    from yolo_dl_b200.synth import HeadSpec
    spec = HeadSpec("thin", 64, [(1030, 3), (2, 1500), (1, 1)],
                    [[(0.1, 0.2), (0.3, 0.1)], [(0.5, 0.4)], [(0.9, 0.8), (0.2, 0.2), (0.4, 0.6)]], 3)
    B = 2
    heads_cpu = synth.make_heads(spec, B, 99)
    heads, detects, merged, infos = build(spec, heads_cpu, dev)
    ref = odet.forward_merged(heads_cpu, infos, spec.num_classes)
    for k in ("cy", "cx", "h", "w"):
        close(getattr(merged, k), ref[k], what=f"thin merged {k}")
    for l, hc in enumerate(heads_cpu):
        out = detects[l].forward(heads[l])
        r = odet.detect2d_forward(hc, spec.num_classes, spec.anchors[l])
        for k in ("cy", "cx", "h", "w"):
            close(getattr(out, k), r[k], what=f"thin layer {l} {k}")
