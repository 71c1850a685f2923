#!/usr/bin/env python
"""Step time against the OFFSET of the gradient tensors inside one arena (everything else fixed):
which address bits does the slow mode depend on?   python tools/placement2.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from yolo_dl_b200 import _lib, synth
from yolo_dl_b200.loss import YoloLossInit, pack_labels
from yolo_dl_b200.pipeline import DeviceStep


def main():
    dev = torch.device("cuda:0")
    spec = synth.yolov4_608()
    B = 64
    gt, off = pack_labels(synth.make_targets(spec, B, 20261105))
    gt, off = gt.to(dev), off.to(dev)
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 20261105)]
    step = DeviceStep(spec, B, dev, loss=YoloLossInit(box_metric="ciou"))
    shapes = [spec.head_shape(l, B) for l in range(step.L)]
    numels = [int(torch.Size(s).numel()) for s in shapes]
    total = sum((n * 4 + 255) // 256 * 256 for n in numels)
    slack = 256 << 20
    arena = torch.empty(total + slack, dtype=torch.uint8, device=dev)
    print(f"heads@{heads[0].data_ptr():#x} arena@{arena.data_ptr():#x}")
    offsets = [0, 512, 4096, 65536, 1 << 20, 2 << 20, 3 << 20, 4 << 20, 8 << 20, 16 << 20, 32 << 20, 48 << 20,
               64 << 20, 96 << 20, 128 << 20, 192 << 20, 0, 2 << 20]
    for o in offsets:
        views, cur = [], o
        for n, s in zip(numels, shapes):
            views.append(arena[cur:cur + n * 4].view(torch.float32).view(s))
            cur += (n * 4 + 255) // 256 * 256
        step.grads = views
        step._grad_ptrs = _lib.ptr_array(views)
        step.capture(heads, gt, off)
        for _ in range(10):
            step.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(300):
                step.replay()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / 300 * 1e3)
        print(f"offset {o:>10d} ({o / 2**20:8.3f} MiB): {min(ts):7.2f} us   grads@{views[0].data_ptr():#x}", flush=True)


if __name__ == "__main__":
    main()
