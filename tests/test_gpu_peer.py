"""The data-parallel loss exchange as peer-memory stores (csrc/peer.cu, dist.PeerLossExchange):
sum_g loss_g * B_g / sum_g B_g of train/src/train/multi_gpu.rs:188-195 without a collective."""
import os
import socket
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_mailbox_single_rank_in_graph(dev):
    """world = 1: the publish kernel inside a CUDA graph advances its own sequence number."""
    from yolo_dl_b200 import dist as ydist
    ex = ydist.PeerLossExchange(dev, local_batch=6, slots=4)
    losses = torch.tensor([1.5, 0.25, 0.5, 0.75], dtype=torch.float32, device=dev)
    comb, seq = ex.collect()
    assert seq == 0
    g = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream(dev)
    with torch.cuda.graph(g, stream=cap):
        ex.publish(losses, cap.cuda_stream)
    for i in range(7):
        losses.fill_(float(i))
        g.replay()
    torch.cuda.synchronize()
    comb, seq = ex.collect()
    assert seq == 7
    assert comb.tolist() == [6.0, 6.0, 6.0, 6.0]
    ex.close()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from yolo_dl_b200 import dist as ydist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    local_b = [3, 5][rank]
    ex = ydist.PeerLossExchange(dev, local_b, slots=4)
    losses = torch.zeros(4, dtype=torch.float32, device=dev)
    for step in range(1, 6):
        losses.copy_(torch.tensor([1.0, 2.0, 3.0, 4.0]) * (rank + 1) * step)
        ex.publish(losses)
    torch.cuda.synchronize()
    dist.barrier()
    comb, seq = ex.collect()
    ref = ydist.weighted_mean_losses(losses, local_b)
    q.put((rank, seq, comb.tolist(), ref.cpu().tolist()))
    dist.barrier()
    ex.close()
    dist.destroy_process_group()


def test_mailbox_two_ranks_equals_nccl_allreduce():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs of one node")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=300) for _ in procs])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, seq, comb, ref in out:
        assert seq == 5
        # step 5: rank 0 publishes 3*[5,10,15,20], rank 1 5*[10,20,30,40]; / 8
        want = [(3 * 5 * k + 5 * 10 * k) / 8 for k in (1, 2, 3, 4)]
        assert comb == pytest.approx(want, rel=1e-6)
        assert comb == pytest.approx(ref, rel=1e-6)


def test_step_graph_publishes_one_step_behind(dev):
    """DeviceStep.capture(publisher=...): replay i stores the scalars of replay i-1 (under its own
    kernels); one explicit publish after the last replay flushes."""
    import yolo_dl_b200 as yh
    from yolo_dl_b200 import dist as ydist, synth
    from yolo_dl_b200.pipeline import DeviceStep
    spec, B = synth.small_spec(), 3
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 5)]
    gt, off = yh.pack_labels(synth.make_targets(spec, B, 5, n_max=6))
    gt, off = gt.to(dev), off.to(dev)
    step = DeviceStep(spec, B, dev)
    ex = ydist.PeerLossExchange(dev, local_batch=B, slots=4)
    step.capture(heads, gt, off, publisher=ex)
    assert step.kernels_in_graph == step.expected_kernels() + 1
    for _ in range(3):
        step.replay()
    torch.cuda.synchronize()
    first = step.losses4.clone()
    comb, seq = ex.collect()
    assert seq == 3 and torch.allclose(comb, first.cpu())       # same inputs: same scalars every step
    heads[0].mul_(0.5)                                           # new inputs -> new loss
    step.replay()
    torch.cuda.synchronize()
    comb, seq = ex.collect()
    assert seq == 4 and torch.allclose(comb, first.cpu())       # still the previous step's
    ex.publish(step.losses4)
    torch.cuda.synchronize()
    comb, seq = ex.collect()
    assert seq == 5 and torch.allclose(comb, step.losses4.cpu()) and not torch.allclose(comb, first.cpu())
    ex.close()
