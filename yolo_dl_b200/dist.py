"""Multi-GPU plumbing of the hot path.

The path shards by image with NO data-path collective: every image's decode, assignment, loss
terms and NMS groups are independent.  One process per GPU owns a contiguous image range, exactly
like `forward_step`'s `narrow` (train/src/train/multi_gpu.rs:545-574).  The only exchange the
reference performs on this path is the minibatch-weighted mean of the loss scalars
(multi_gpu.rs:188-195, loss_.rs:527-577, tch-goodies/src/tensor.rs:66-80); here it is ONE
all-reduce(sum) of 5 floats over NCCL (gloo in the CPU tests).
"""
from typing import Tuple

import torch
import torch.distributed as dist


def shard_range(batch: int, rank: int, world: int) -> Tuple[int, int]:
    """contiguous [start, end) of images of `rank`; sizes differ by at most one."""
    base, rem = divmod(batch, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def weighted_mean_losses(losses4: torch.Tensor, local_batch: int, group=None) -> torch.Tensor:
    """YoloLossOutput::weighted_mean over the ranks: sum_g loss_g * B_g / sum_g B_g.
    `losses4` = [total, iou, classification, objectness] of this rank's LOCAL means (reference
    behaviour: per-device means, not a global mean).  Returns the 4 combined scalars."""
    buf = torch.empty(5, dtype=torch.float32, device=losses4.device)
    buf[:4] = losses4.detach().float() * float(local_batch)
    buf[4] = float(local_batch)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[:4] / buf[4]


def gradient_scale(local_batch: int, global_batch: int) -> float:
    """multi_gpu.rs:619-623,690-693: each replica's gradient is weighted by its minibatch size
    and the sum divided by the global batch; with an all-reduce(sum) of the caller's backbone
    gradients this is the per-rank factor."""
    return float(local_batch) / float(global_batch)


class LossAllReduce:
    """The per-step loss exchange off the critical path: preallocated 5-float buffer, its own
    stream, so that the (latency-bound, ~30 us) NCCL all-reduce of step i runs under the kernels
    of step i+1.  `launch(losses4)` enqueues; `result()` returns the combined [total, iou, cls,
    obj] of the last launch (synchronises that stream only)."""

    def __init__(self, device: torch.device, local_batch: int, group=None):
        self.group = group
        self.on_cuda = device.type == "cuda"
        self.src = torch.zeros(5, dtype=torch.float32, device=device)
        self.buf = torch.zeros(5, dtype=torch.float32, device=device)
        self.local_batch = float(local_batch)
        self.src[4] = self.local_batch
        self.launched = False
        self.n_launched = 0
        if self.on_cuda:
            self.ready2 = [torch.cuda.Event(), torch.cuda.Event()]
            self.done2 = [torch.cuda.Event(), torch.cuda.Event()]
        self.active = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        if self.on_cuda:
            self.stream = torch.cuda.Stream(device)
            self.taken = torch.cuda.Event()
            self.done = torch.cuda.Event()

    def _enqueue(self, losses4):
        torch.mul(losses4.detach(), self.local_batch, out=self.src[:4])
        self.buf.copy_(self.src)

    def launch(self, losses4: torch.Tensor, wait_taken: bool = True):
        """`wait_taken=False`: the caller guarantees that `losses4` is not overwritten before
        the NEXT launch (DeviceStep.capture(loss_slots=2) alternates two buffers), so the
        producing stream is not made to wait for this launch's read — only for the previous one."""
        if not self.on_cuda:
            self._enqueue(losses4)
            if self.active:
                dist.all_reduce(self.buf, op=dist.ReduceOp.SUM, group=self.group)
            return
        cur = torch.cuda.current_stream(losses4.device)
        if not wait_taken and self.launched:
            cur.wait_event(self.taken)  # the previous launch has read its buffer
        self.launched = True
        ready = torch.cuda.Event()
        ready.record(cur)
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            self._enqueue(losses4)
            self.taken.record(self.stream)
            if self.active:
                dist.all_reduce(self.buf, op=dist.ReduceOp.SUM, group=self.group)
            self.done.record(self.stream)
        if wait_taken:
            cur.wait_event(self.taken)  # losses4 may be overwritten by the next step from here on

    def result(self) -> torch.Tensor:
        if self.on_cuda:
            self.done.synchronize()
        return self.buf[:4] / self.buf[4]

    def launch_weighted(self, weighted5: torch.Tensor):
        """`weighted5` = [loss*B_local (4), B_local] already on the device (DeviceStep.weighted5,
        written inside the step graph into one of two alternating buffers): the all-reduce runs in
        place on it, on this object's stream; nothing else is launched and the producing stream
        never waits for it (only for the launch before the previous one)."""
        cur = torch.cuda.current_stream(weighted5.device)
        k = self.n_launched & 1
        self.n_launched += 1
        if self.n_launched > 2:
            cur.wait_event(self.done2[k])  # same buffer, two launches ago: finished long since
        self.ready2[k].record(cur)
        self.stream.wait_event(self.ready2[k])
        if self.active:
            with torch.cuda.stream(self.stream):
                dist.all_reduce(weighted5, op=dist.ReduceOp.SUM, group=self.group)
        self.done2[k].record(self.stream)
        self.buf = weighted5
        self.done = self.done2[k]


class PeerLossExchange:
    """The same exchange WITHOUT a collective: every rank's step writes its five weighted scalars
    straight into all peers' mailboxes over NVLink / NVSwitch (CUDA IPC peer memory) from a
    32-thread kernel that is part of the step's CUDA graph — no NCCL kernel, no host launch per
    step (include/yolohot.h §8e, csrc/peer.cu).  `publish(losses4, stream)` enqueues (capturable);
    `collect()` returns (combined [total, iou, cls, obj], sequence number) of the newest step all
    ranks have published.  One node only (CUDA IPC); torch.distributed is used once, to exchange
    the 64-byte handles."""

    def __init__(self, device: torch.device, local_batch: int, slots: int = 8, group=None):
        import ctypes
        from . import _lib
        self._lib_mod, self.lib = _lib, _lib.load()
        self.dev, self.weight = device, float(local_batch)
        active = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if active else 1
        self.rank = dist.get_rank(group) if active else 0
        self.handle = ctypes.c_void_p()
        mail = ctypes.c_void_p()
        h = (ctypes.c_ubyte * 64)()
        _lib.check(self.lib.yh_peer_mailbox_create(device.index or 0, self.world, slots,
                                                   ctypes.byref(self.handle), ctypes.byref(mail), h))
        mine = torch.tensor(list(h), dtype=torch.uint8, device=device)
        if self.world > 1:
            allh = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(allh, mine, group=group)
        else:
            allh = [mine]
        self.peers = (ctypes.c_void_p * self.world)()
        self._opened = []
        for r in range(self.world):
            if r == self.rank:
                self.peers[r] = mail.value
                continue
            hb = (ctypes.c_ubyte * 64)(*allh[r].cpu().tolist())
            p = ctypes.c_void_p()
            _lib.check(self.lib.yh_peer_mailbox_open(device.index or 0, hb, ctypes.byref(p)))
            self.peers[r] = p.value
            self._opened.append(p.value)
        if self.world > 1:
            dist.barrier(group=group)  # every mailbox is mapped before anyone publishes

    def publish(self, losses4: torch.Tensor, stream=None):
        s = stream if stream is not None else torch.cuda.current_stream(self.dev).cuda_stream
        self._lib_mod.check(self.lib.yh_loss_publish(self.handle, losses4.data_ptr(), self.weight,
                                                     self.peers, self.rank, s))

    def collect(self):
        import ctypes
        out = (ctypes.c_float * 4)()
        seq = ctypes.c_uint32(0)
        self._lib_mod.check(self.lib.yh_loss_collect(self.handle, out, ctypes.byref(seq),
                                                     torch.cuda.current_stream(self.dev).cuda_stream))
        return torch.tensor(list(out), dtype=torch.float32), int(seq.value)

    def close(self):
        for p in self._opened:
            self.lib.yh_peer_mailbox_close(self.dev.index or 0, p)
        self._opened = []
        if self.handle:
            self.lib.yh_peer_mailbox_destroy(self.handle)
            self.handle = None
