"""World-size-2 gloo test of the data-parallel gradient path (yolo_dl_b200/dp.py):
the all-reduce form equals the reference's master-device aggregation
sum_r(grad_r * B_r) / B (train/src/train/multi_gpu.rs:619-623,671-693)."""
import os
import socket
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_params(seed):
    g = torch.Generator().manual_seed(seed)
    shapes = [(16, 3, 3, 3), (16,), (32, 16, 1, 1), (5,), (7, 11)]
    return [torch.randn(s, generator=g).requires_grad_(True) for s in shapes]


def _worker(rank, world, port, q, exact):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from yolo_dl_b200 import dp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    params = _make_params(100 + rank)          # ranks start with DIFFERENT weights
    dp.sync_weights(params, src=0)
    w0 = [p.detach().clone() for p in params]
    local_b = [3, 5][rank]
    gar = dp.GradientAllReduce(params, bucket_bytes=2048)   # small buckets: several collectives
    g = torch.Generator().manual_seed(7 + rank)
    for p in params:
        p.grad.copy_(torch.randn(p.shape, generator=g))
    gar.reduce(local_b, 8, exact_order=exact)
    q.put((rank, [w.tolist() for w in w0], [p.grad.tolist() for p in params], len(gar.buckets)))
    dist.destroy_process_group()


def _run(exact):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, exact)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=120) for _ in procs])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


def _expected():
    grads = []
    for rank in range(2):
        g = torch.Generator().manual_seed(7 + rank)
        grads.append([torch.randn(p.shape, generator=g) for p in _make_params(0)])
    return grads


def test_gradient_all_reduce_exact_order_gloo_world2():
    out = _run(True)
    (r0, w0, g0, nb0), (r1, w1, g1, nb1) = out
    assert w0 == w1 == [p.tolist() for p in _make_params(100)]     # sync_weights: rank 0's weights
    assert nb0 > 1
    grads = _expected()
    for k in range(len(g0)):
        want = (grads[0][k] * 3.0 + grads[1][k] * 5.0) / 8.0        # multi_gpu.rs:619-623,690-693
        assert torch.equal(torch.tensor(g0[k]), want) and torch.equal(torch.tensor(g1[k]), want)


def test_gradient_all_reduce_folded_factor_gloo_world2():
    out = _run(False)
    (_, _, g0, _), (_, _, g1, _) = out
    grads = _expected()
    for k in range(len(g0)):
        want = (grads[0][k] * 3.0 + grads[1][k] * 5.0) / 8.0
        assert g0[k] == g1[k]
        torch.testing.assert_close(torch.tensor(g0[k]), want, rtol=1e-6, atol=1e-7)


def test_single_process_is_a_local_scale():
    sys.path.insert(0, ROOT)
    from yolo_dl_b200 import dp
    params = _make_params(1)
    gar = dp.GradientAllReduce(params)
    for p in params:
        p.grad.fill_(2.0)
    gar.reduce(4, 8)
    assert all(bool((p.grad == 1.0).all()) for p in params)
    gar.zero_grad()
    assert all(bool((p.grad == 0.0).all()) for p in params)


def test_reduce_repairs_grads_dropped_by_set_to_none():
    """optimizer.zero_grad(set_to_none=True) drops the bucket views; the next backward allocates
    fresh .grad tensors.  reduce() must pick those up (and re-point .grad at the buckets) instead
    of averaging stale zero buckets."""
    sys.path.insert(0, ROOT)
    from yolo_dl_b200 import dp
    params = _make_params(2)
    gar = dp.GradientAllReduce(params)
    opt = torch.optim.SGD(params, lr=0.1)
    opt.zero_grad(set_to_none=True)
    assert all(p.grad is None for p in params)
    sum((p * p).sum() for p in params[:-1]).backward()     # the last parameter gets no gradient
    want = [2.0 * p.detach() * 0.5 for p in params[:-1]]
    gar.reduce(4, 8)
    for p, w in zip(params[:-1], want):
        assert torch.equal(p.grad, w)
    assert bool((params[-1].grad == 0).all())
    ptrs = {b.data_ptr() for b in gar.buckets}
    assert all(p.grad.data_ptr() in range(min(ptrs), max(ptrs) + gar.nbytes()) for p in params)
    gar.zero_grad()
    sum((p * p).sum() for p in params).backward()
    assert torch.equal(gar._views[0], params[0].grad)      # accumulates straight into the bucket
