"""Two devices in ONE process, one OS thread per device — the reference's multi-GPU calling
convention (train/src/train/multi_gpu.rs:578-646, `par_map_unordered` over per-device workers) —
and the data-parallel gradient all-reduce over NCCL on real GPUs.
Both need >= 2 GPUs: they skip on the 1-GPU test box and are run with `gpurun --gpus 2`
(log kept in profiles/)."""
import os
import socket
import sys
import threading

import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh
from yolo_dl_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker_step(device_index, heads_cpu, targets, spec, out):
    try:
        dev = torch.device("cuda", device_index)
        torch.cuda.set_device(dev)
        heads = [h.to(dev).requires_grad_(True) for h in heads_cpu]
        detects = [yh.Detect2DInit(spec.num_classes, a).build() for a in spec.anchors]
        merged = yh.MergedDenseDetection.from_heads(heads, detects)
        loss = yh.YoloLossInit(box_metric="ciou").build()
        res = []
        for _ in range(3):   # several rounds so that the two threads really overlap
            for h in heads:
                h.grad = None
            o, aux = loss.forward(merged, targets)
            o.total_loss.backward()
            # capacity above YH_BIG_MIN: the large-segment kernel (cooperative launch, > 48 KB dynamic
            # shared memory, a per-DEVICE function attribute) is launched on this device too
            nms = yh.NonMaxSuppressionInit(0.5, 0.25, False, max_candidates_per_image=16384).build()
            kept = nms.forward(merged)
            ltrb, sc, grp = synth.make_nms_boxes(9000, 3, num_groups=1)
            big = yh.nms_by_scores(torch.from_numpy(ltrb).to(dev), torch.from_numpy(sc).to(dev),
                                   torch.from_numpy(grp).to(dev), 0.5)
            res.append((float(o.total_loss), aux.matchings.pred_indexes.flats.cpu().numpy(),
                        [h.grad.cpu() for h in heads], kept.instances.cpu().numpy(), big.cpu().numpy()))
        torch.cuda.synchronize(dev)
        out[device_index] = res
    except Exception as e:   # surfaced by the main thread
        out[device_index] = e


def test_two_threads_two_devices_one_process():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    spec = synth.small_spec()
    B = 4
    heads_cpu = synth.make_heads(spec, B, 21, obj_mean=0.5, cls_mean=0.0)
    targets = synth.make_targets(spec, B, 21, n_max=9)
    out = {}
    # device 1 FIRST, then both at once: a process-wide "attribute already set" flag would break
    # whichever device comes second
    _worker_step(1, heads_cpu, targets, spec, out)
    assert not isinstance(out[1], Exception), out[1]
    first = out.pop(1)
    ts = [threading.Thread(target=_worker_step, args=(d, heads_cpu, targets, spec, out)) for d in (0, 1)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    for d in (0, 1):
        assert not isinstance(out[d], Exception), (d, out[d])
        for r in out[d]:
            ref = first[0]
            assert r[0] == ref[0]
            assert np.array_equal(r[1], ref[1]) and np.array_equal(r[3], ref[3]) and np.array_equal(r[4], ref[4])
            for a, b in zip(r[2], ref[2]):
                assert torch.equal(a, b)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from yolo_dl_b200 import dp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    g = torch.Generator().manual_seed(100 + rank)
    shapes = [(64, 32, 3, 3), (64,), (256, 64, 1, 1), (5,), (255, 256, 1, 1)]
    params = [torch.randn(s, generator=g).to(dev).requires_grad_(True) for s in shapes]
    dp.sync_weights(params, src=0)
    w0 = [p.detach().cpu() for p in params]
    local_b = [3, 5][rank]
    res = {}
    for exact in (True, False):
        gar = dp.GradientAllReduce(params, bucket_bytes=64 << 10)
        gg = torch.Generator().manual_seed(7 + rank)
        for p in params:
            p.grad.copy_(torch.randn(p.shape, generator=gg))
        gar.reduce(local_b, 8, exact_order=exact)
        torch.cuda.synchronize()
        res[exact] = ([p.grad.cpu() for p in params], len(gar.buckets))
    q.put((rank, w0, res))
    dist.destroy_process_group()


def test_gradient_all_reduce_nccl_two_gpus():
    """f3 on hardware: the bucketed NCCL all-reduce equals the reference's master-device
    aggregation sum_g(grad_g * B_g) / B (multi_gpu.rs:619-623,671-693) — bit for bit in the
    exact-order form, to rounding with the folded pre-multiplier."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=300) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    (_, w0, r0), (_, w1, r1) = out
    for a, b in zip(w0, w1):
        assert torch.equal(a, b)          # sync_weights: rank 0's weights everywhere
    shapes = [tuple(t.shape) for t in w0]
    grads = []
    for rank in range(2):
        gg = torch.Generator().manual_seed(7 + rank)
        grads.append([torch.randn(s, generator=gg) for s in shapes])
    for k in range(len(shapes)):
        want = (grads[0][k] * 3.0 + grads[1][k] * 5.0) / 8.0
        assert torch.equal(r0[True][0][k], want) and torch.equal(r1[True][0][k], want)
        assert torch.equal(r0[False][0][k], r1[False][0][k])
        torch.testing.assert_close(r0[False][0][k], want, rtol=1e-6, atol=1e-7)
    assert r0[True][1] > 1
