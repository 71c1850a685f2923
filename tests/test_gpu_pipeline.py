"""DeviceStep / HostStep: the whole step on preallocated buffers.  The sequential form, the
fork/join form over four streams and its CUDA-graph replay must give bit-identical results
(every reduction is fixed-order), and all of them must agree with the oracle."""
import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh
from oracle import detect as odet, loss as oloss, nms as onms
from yolo_dl_b200 import synth
from yolo_dl_b200.pipeline import DeviceStep, HostStep
from util import close, grad_close

pytestmark = pytest.mark.gpu


def snapshot(step):
    torch.cuda.synchronize()
    M, K = step.counts()
    return {
        "M": M, "K": K,
        "losses": step.losses4.clone(),
        "boxes": step.boxes.clone(),
        "m_batch": step.m_batch[:M].clone(), "m_flat": step.m_flat[:M].clone(),
        "m_gt": step.m_gt[:M].clone(), "iou": step.iou_score[:M].clone(),
        "kb": step.kb[:K].clone(), "kc": step.kc[:K].clone(), "ki": step.ki[:K].clone(),
        "tlbr": step.tlbr[:K].clone(), "conf": step.conf[:K].clone(),
        "grads": [g.clone() for g in step.grads],
    }


def same(a, b):
    assert a["M"] == b["M"] and a["K"] == b["K"]
    for k in a:
        if k in ("M", "K"):
            continue
        xs, ys = (a[k], b[k]) if isinstance(a[k], list) else ([a[k]], [b[k]])
        for x, y in zip(xs, ys):
            assert torch.equal(x, y), k


@pytest.mark.parametrize("name,B", [("small", 3), ("yolov4_608", 4), ("yolov4_tiny_416", 16)])
def test_overlapped_and_graph_equal_sequential_and_oracle(dev, name, B):
    spec = synth.small_spec() if name == "small" else getattr(synth, name)()
    om_, cm_ = (-2.0, -1.0) if name == "small" else (-3.5, -2.5)
    heads_cpu = synth.make_heads(spec, B, synth.SEED0 + 5, obj_mean=om_, cls_mean=cm_)
    targets = synth.make_targets(spec, B, synth.SEED0 + 5, n_max=12)
    heads = [h.to(dev) for h in heads_cpu]
    gt_h, off_h = yh.pack_labels(targets)
    gt, off = gt_h.to(dev), off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=0.3, nms_iou=0.5, by_class=True, cand_per_image=16384,
                      max_detections=B * 16384)

    step.run(heads, gt, off)
    seq = snapshot(step)

    for g in step.grads:
        g.fill_(float("nan"))
    step.losses4.fill_(float("nan"))
    step.run_overlapped(heads, gt, off)
    same(seq, snapshot(step))

    step.capture(heads, gt, off)
    assert step.kernels_in_graph == step.expected_kernels()
    for _ in range(3):
        for g in step.grads:
            g.fill_(float("nan"))
        step.losses4.fill_(float("nan"))
        step.boxes.fill_(float("nan"))
        step.replay()
        same(seq, snapshot(step))

    # and against the oracle
    infos = odet.make_infos(spec.shapes, spec.anchors)
    ref = oloss.yolo_loss(heads_cpu, infos, spec.num_classes, targets, oloss.LossConfig(box_metric="ciou"))
    close(seq["losses"][0], ref["losses"]["total"], what="total")
    assert np.array_equal(seq["m_flat"].cpu().numpy(), ref["matches"][1])
    for g, r in zip(seq["grads"], ref["grads"]):
        grad_close(g, r, what="grad")
    om = odet.forward_merged(heads_cpu, infos, spec.num_classes)
    close(seq["boxes"][0], om["cy"], what="cy")
    ok = onms.nms_forward(om, 0.3, 0.5, True)
    assert np.array_equal(seq["ki"].cpu().numpy(), ok["instances"])
    assert np.array_equal(seq["kc"].cpu().numpy(), ok["classes"])


@pytest.mark.parametrize("name,B,group", [("small", 12, 4), ("small", 8, 8), ("yolov4_608", 16, 8),
                                          ("yolov4_tiny_416", 20, 8)])
def test_grouped_schedule_equals_sequential(dev, name, B, group, monkeypatch):
    """The loss in L2-sized image ranges (yh_loss_dense_range -> yh_loss_rows_range per range,
    gradients of the matched rows stored straight over the dense gradient, then
    yh_loss_finish_ranges): every output is bit-identical to the one-sequence form except the four
    loss scalars, whose f64 partial sums are associated differently (equal to f32 rounding).
    Covers a last range shorter than the others and the graph replay."""
    monkeypatch.setenv("YH_GROUP", str(group))
    spec = synth.small_spec() if name == "small" else getattr(synth, name)()
    om_, cm_ = (-2.0, -1.0) if name == "small" else (-3.5, -2.5)
    heads = [h.to(dev) for h in synth.make_heads(spec, B, synth.SEED0 + 6, obj_mean=om_, cls_mean=cm_)]
    targets = synth.make_targets(spec, B, synth.SEED0 + 6, n_max=12, clustered=0.3)
    targets[1] = []                      # an image without labels inside a range
    gt_h, off_h = yh.pack_labels(targets)
    gt, off = gt_h.to(dev), off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=0.3, nms_iou=0.5, by_class=False, cand_per_image=4096,
                      max_detections=B * 4096)
    step.run(heads, gt, off)
    seq = snapshot(step)

    def check():
        got = snapshot(step)
        torch.testing.assert_close(got["losses"], seq["losses"], rtol=2e-7, atol=0)
        got["losses"] = seq["losses"]
        same(seq, got)

    for g in step.grads:
        g.fill_(float("nan"))
    step.losses4.fill_(float("nan"))
    step.run_grouped(heads, gt, off)
    check()
    step.schedule = "grouped"
    step.capture(heads, gt, off)
    assert step.kernels_in_graph == step.expected_kernels(True)
    for _ in range(2):
        for g in step.grads:
            g.fill_(float("nan"))
        step.losses4.fill_(float("nan"))
        step.replay()
        check()
    # without the per-image offsets of yh_assign the rows kernel searches the batch column itself
    R = (B + group - 1) // group
    hp = yh._lib.ptr_array(heads)
    s = torch.cuda.current_stream(dev).cuda_stream
    for g in step.grads:
        g.fill_(float("nan"))
    for r in range(R):
        lo, hi = r * group, min(B, (r + 1) * group)
        yh._lib.check(step.lib.yh_loss_dense_range(step.ws.handle, hp, step.layers, step.L, B, step.C, lo, hi,
                                                   r, R, step.cfg, step._grad_ptrs, s))
        yh._lib.check(step.lib.yh_loss_rows_range(
            step.ws.handle, hp, step.layers, step.L, B, step.C, lo, hi, r, R, step.m_batch.data_ptr(),
            step.m_flat.data_ptr(), step.m_target.data_ptr(), step.m_class.data_ptr(), step.M.data_ptr(),
            step.mcap, None, step.cfg, step.iou_score.data_ptr(), step.pred_rows.data_ptr(),
            step._grad_ptrs, s))
    yh._lib.check(step.lib.yh_loss_finish_ranges(step.ws.handle, step.layers, step.L, B, step.C, R,
                                                 step.M.data_ptr(), step.mcap, step.cfg,
                                                 step.losses4.data_ptr(), s))
    check()
    # an image range must start on a 16-byte boundary of every head tensor
    if name == "small":
        rc = step.lib.yh_loss_dense_range(step.ws.handle, hp, step.layers, step.L, B, step.C, 1, 4, 0, 2,
                                          step.cfg, step._grad_ptrs, s)
        assert rc == yh._lib.YH_ERR_INVALID and b"multiple of 4" in step.lib.yh_last_error()


@pytest.mark.parametrize("name,B", [("small", 5), ("yolov4_608", 6), ("yolov4_tiny_416", 9)])
def test_layered_schedule_equals_sequential(dev, name, B):
    """The loss cut by layer (yh_loss_dense_layers per layer, yh_loss_finish_layers of a layer right
    behind its dense pass, the last one reducing): every output bit-identical to the one-sequence
    form except the four loss scalars (differently associated f64 partial sums)."""
    spec = synth.small_spec() if name == "small" else getattr(synth, name)()
    om_, cm_ = (-2.0, -1.0) if name == "small" else (-3.5, -2.5)
    heads = [h.to(dev) for h in synth.make_heads(spec, B, synth.SEED0 + 7, obj_mean=om_, cls_mean=cm_)]
    targets = synth.make_targets(spec, B, synth.SEED0 + 7, n_max=12, clustered=0.3)
    targets[0] = []
    gt_h, off_h = yh.pack_labels(targets)
    gt, off = gt_h.to(dev), off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=0.3, nms_iou=0.5, by_class=False, cand_per_image=4096,
                      max_detections=B * 4096)
    step.run(heads, gt, off)
    seq = snapshot(step)

    def check():
        got = snapshot(step)
        torch.testing.assert_close(got["losses"], seq["losses"], rtol=2e-7, atol=0)
        got["losses"] = seq["losses"]
        same(seq, got)

    for g in step.grads:
        g.fill_(float("nan"))
    step.losses4.fill_(float("nan"))
    step.run_layered(heads, gt, off)
    check()
    step.schedule = "layered"
    step.capture(heads, gt, off)
    assert step.kernels_in_graph == step.expected_kernels()
    for _ in range(2):
        for g in step.grads:
            g.fill_(float("nan"))
        step.losses4.fill_(float("nan"))
        step.replay()
        check()


def test_finish_after_dense_equals_finish(dev, monkeypatch):
    """yh_loss_finish_after_dense (the scatter launched with PDL behind the dense pass, loads before
    the wait) gives the bits of yh_loss_finish, eagerly and in a replayed graph on refilled inputs."""
    monkeypatch.setenv("YH_FINISH_EARLY", "1")
    spec = synth.yolov4_tiny_416()
    B = 20
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 31, obj_mean=-3.5, cls_mean=-2.5)]
    gt_h, off_h = yh.pack_labels(synth.make_targets(spec, B, 31, n_max=12, clustered=0.3))
    G = 512
    gt = torch.zeros((G, 5), dtype=torch.float64, device=dev)
    gt[:gt_h.shape[0]] = gt_h.to(dev)
    off = off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=0.3, cand_per_image=4096, max_detections=B * 4096)
    step.run(heads, gt, off)
    seq = snapshot(step)
    step.run_overlapped(heads, gt, off)
    same(seq, snapshot(step))
    step.capture(heads, gt, off)
    # new labels and heads in the same buffers
    for d, h in zip(heads, synth.make_heads(spec, B, 32, obj_mean=-3.5, cls_mean=-2.5)):
        d.copy_(h)
    gt2, off2 = yh.pack_labels(synth.make_targets(spec, B, 32, n_max=12))
    gt[:gt2.shape[0]] = gt2.to(dev)
    off.copy_(off2)
    monkeypatch.setenv("YH_FINISH_EARLY", "0")
    step.run(heads, gt, off)
    seq2 = snapshot(step)
    for g in step.grads:
        g.fill_(float("nan"))
    step.replay()
    same(seq2, snapshot(step))


def test_graph_replay_follows_refilled_inputs(dev):
    spec = synth.small_spec()
    B = 4
    step = DeviceStep(spec, B, dev, nms_conf=0.3)
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 1, obj_mean=-1.0, cls_mean=-1.0)]
    t1 = synth.make_targets(spec, B, 1, n_max=5)
    gt_h, off_h = yh.pack_labels(t1)
    G = 64
    gt = torch.zeros((G, 5), dtype=torch.float64, device=dev)
    off = off_h.to(dev)
    gt[:gt_h.shape[0]] = gt_h.to(dev)
    step.capture(heads, gt, off)
    # new batch, same buffers
    heads2 = synth.make_heads(spec, B, 2, obj_mean=-1.0, cls_mean=-1.0)
    t2 = synth.make_targets(spec, B, 2, n_max=5)
    gt2, off2 = yh.pack_labels(t2)
    for d, h in zip(heads, heads2):
        d.copy_(h)
    gt[:gt2.shape[0]] = gt2.to(dev)
    off.copy_(off2)
    step.replay()
    got = snapshot(step)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    ref = oloss.yolo_loss(heads2, infos, spec.num_classes, t2, oloss.LossConfig(box_metric="ciou"))
    close(got["losses"][0], ref["losses"]["total"], what="total")
    assert np.array_equal(got["m_flat"].cpu().numpy(), ref["matches"][1])


def test_host_step_round_trip(dev):
    spec = synth.small_spec()
    B = 5
    heads_cpu = synth.make_heads(spec, B, 9, obj_mean=-1.0, cls_mean=-1.0)
    targets = synth.make_targets(spec, B, 9, n_max=6)
    gt_h, off_h = yh.pack_labels(targets)
    hs = HostStep(spec, B, dev, nms_conf=0.3)
    pinned = [h.pin_memory() for h in heads_cpu]
    for _ in range(2):
        losses, M, K = hs.run(pinned, gt_h.pin_memory(), off_h.pin_memory())
    infos = odet.make_infos(spec.shapes, spec.anchors)
    ref = oloss.yolo_loss(heads_cpu, infos, spec.num_classes, targets, oloss.LossConfig(box_metric="ciou"))
    close(losses[0], ref["losses"]["total"], what="total")
    assert M == len(ref["matches"][1])
    assert np.array_equal(hs.h_match[1][:M].numpy(), ref["matches"][1])
    for g, r in zip(hs.h_grads, ref["grads"]):
        grad_close(g, r, what="grad")
    assert hs.h2d_bytes > 0 and hs.d2h_bytes > 0


def test_host_pipeline_overlapped_slots(dev):
    """Results of the 2-slot pipeline equal the unpipelined HostStep for a stream of different batches."""
    from yolo_dl_b200.pipeline import HostPipeline
    spec = synth.small_spec()
    B = 4
    batches = []
    for seed in range(5):
        heads = [h.pin_memory() for h in synth.make_heads(spec, B, 100 + seed, obj_mean=-1.0, cls_mean=-1.0)]
        gt, off = yh.pack_labels(synth.make_targets(spec, B, 100 + seed, n_max=6))
        batches.append((heads, gt, off))
    hs = HostStep(spec, B, dev, nms_conf=0.3)
    want = []
    for heads, gt, off in batches:
        losses, M, K = hs.run(heads, gt, off)
        want.append((losses.clone(), M, K, hs.h_match[:, :M].clone(), hs.h_det[:, :K].clone(),
                     [g.clone() for g in hs.h_grads]))
    pipe = HostPipeline(spec, B, dev, depth=2, nms_conf=0.3)
    slots = []
    got = []

    def collect(k):
        losses, M, K = pipe.result(k)
        s = pipe.slots[k]
        got.append((losses.clone(), M, K, s.h_match[:, :M].clone(), s.h_det[:, :K].clone(),
                    [g.clone() for g in s.h_grads]))

    for i, (heads, gt, off) in enumerate(batches):
        if i >= 2:
            collect(slots[i - 2])  # a slot's results must be read before it is reused
        slots.append(pipe.submit(heads, gt, off))
    collect(slots[-2])
    collect(slots[-1])
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert torch.equal(g[0], w[0]) and g[1] == w[1] and g[2] == w[2]
        assert torch.equal(g[3], w[3]) and torch.equal(g[4], w[4])
        for x, y in zip(g[5], w[5]):
            assert torch.equal(x, y)


def test_capture_with_two_loss_slots_alternates_buffers(dev):
    spec = synth.small_spec()
    B = 3
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 4, obj_mean=-1.0, cls_mean=-1.0)]
    gt_h, off_h = yh.pack_labels(synth.make_targets(spec, B, 4, n_max=5))
    gt, off = gt_h.to(dev), off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=0.3)
    step.run(heads, gt, off)
    torch.cuda.synchronize()
    want = step.losses4.clone()
    step.capture(heads, gt, off, loss_slots=2)
    seen = []
    for _ in range(4):
        step.losses4.fill_(float("nan"))
        step.replay()
        torch.cuda.synchronize()
        assert torch.equal(step.losses4, want)
        seen.append(step.losses4.data_ptr())
    assert seen[0] == seen[2] and seen[1] == seen[3] and seen[0] != seen[1]
    assert torch.equal(step.weighted5[:4], want * B) and float(step.weighted5[4]) == B
