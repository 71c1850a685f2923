"""Data-parallel gradient path either side of the hot path.

The reference trains one replica per GPU inside ONE process and moves everything through GPU 0:
`sync_weights` copies every variable of worker 0 to the other workers each step
(train/src/train/multi_gpu.rs:503-530) and `backward_step` brings every worker's gradients
(pre-multiplied by its minibatch size, :619-623) to the master device with `to_device`, adds them
up, divides by the batch size (:671-693) and steps the master optimizer only (:697-722) — two
|params| x (G-1) P2P transfers through one GPU per step.

One process per GPU over NCCL replaces both: every rank holds the same weights, contributes
`grad * B_local`, one all-reduce(sum) over NVLink/NVSwitch per bucket, `/ B_global`, and every
rank takes the same optimizer step, so weights never need re-broadcasting after the initial
`sync_weights`.  Gradients live in a few large flat buckets (the parameters' `.grad` are views
into them), so a step is a handful of collectives sized for launch latency, not one per tensor.
"""
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist


def _world(group=None) -> int:
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def sync_weights(params: Iterable[torch.Tensor], src: int = 0, group=None) -> None:
    """multi_gpu.rs:503-530 as a broadcast from `src` (call once after construction / restore)."""
    if _world(group) == 1:
        return
    with torch.no_grad():
        for p in params:
            dist.broadcast(p.data, src=src, group=group)


class GradientAllReduce:
    """Bucketed weighted-mean all-reduce of parameter gradients.

        gar = GradientAllReduce(model_parameters)        # .grad become views into flat buckets
        loss.backward()
        gar.reduce(local_batch, global_batch)            # grad <- sum_r grad_r * B_r / B
        optimizer.step()

    `exact_order=True` keeps the reference's three roundings (multiply by B_local, sum, divide by
    B, multi_gpu.rs:619-623,690-693); the default folds both scalings into the collective's
    pre-multiplier B_local/B (NCCL PreMulSum: no extra pass over the gradients; other backends:
    one multiply), which differs from the reference by rounding only (<= 2 ulp per element).
    """

    def __init__(self, params: Iterable[torch.Tensor], bucket_bytes: int = 256 << 20, group=None):
        self.group = group
        self.params: List[torch.Tensor] = [p for p in params if p.requires_grad]
        self.buckets: List[torch.Tensor] = []
        cur, cur_bytes, key = [], 0, None
        plan = []
        for p in self.params:
            k = (p.device, p.dtype)
            nbytes = p.numel() * p.element_size()
            if cur and (k != key or cur_bytes + nbytes > bucket_bytes):
                plan.append(cur)
                cur, cur_bytes = [], 0
            key = k
            cur.append(p)
            cur_bytes += nbytes
        if cur:
            plan.append(cur)
        self._views: List[torch.Tensor] = []   # per parameter, its slice of a bucket
        for ps in plan:
            flat = torch.zeros(sum(p.numel() for p in ps), dtype=ps[0].dtype, device=ps[0].device)
            o = 0
            for p in ps:
                v = flat[o:o + p.numel()].view_as(p)
                if p.grad is not None:
                    v.copy_(p.grad)
                p.grad = v
                self._views.append(v)
                o += p.numel()
            self.buckets.append(flat)

    def zero_grad(self) -> None:
        """Zeroes the buckets and keeps `.grad` pointing into them.  Use this instead of
        `optimizer.zero_grad()` / `model.zero_grad()`, whose default `set_to_none=True` drops the
        views (reduce() then repairs the aliasing, at the price of one copy per parameter)."""
        for b in self.buckets:
            b.zero_()
        self._reattach(copy=False)

    def _reattach(self, copy: bool) -> None:
        """Every `.grad` must alias its bucket slice, or the all-reduce averages stale buckets
        while autograd accumulates elsewhere (after zero_grad(set_to_none=True) the next
        backward allocates fresh .grad tensors).  A detached gradient is copied into its slice
        and re-pointed; a missing one (parameter unused this step) counts as zero."""
        for p, v in zip(self.params, self._views):
            g = p.grad
            if g is not None and g.data_ptr() == v.data_ptr() and g.shape == v.shape:
                continue
            if copy:
                if g is None:
                    v.zero_()
                else:
                    v.copy_(g)
            p.grad = v

    def reduce(self, local_batch: int, global_batch: int, exact_order: bool = False, async_op: bool = False):
        """In place on the buckets.  Returns the list of work handles when `async_op`."""
        world = _world(self.group)
        handles = []
        with torch.no_grad():
            self._reattach(copy=True)
            for b in self.buckets:
                if exact_order:
                    b.mul_(float(local_batch))
                    if world > 1:
                        dist.all_reduce(b, op=dist.ReduceOp.SUM, group=self.group)
                    b.div_(float(global_batch))
                    continue
                factor = float(local_batch) / float(global_batch)
                if world > 1 and b.is_cuda and dist.get_backend(self.group) == "nccl" and \
                        hasattr(dist, "_make_nccl_premul_sum"):
                    h = dist.all_reduce(b, op=dist._make_nccl_premul_sum(factor), group=self.group,
                                        async_op=async_op)
                else:
                    b.mul_(factor)
                    h = dist.all_reduce(b, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op) \
                        if world > 1 else None
                if async_op and h is not None:
                    handles.append(h)
        return handles

    def nbytes(self) -> int:
        return sum(b.numel() * b.element_size() for b in self.buckets)
