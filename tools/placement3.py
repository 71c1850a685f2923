#!/usr/bin/env python
"""Re-capturing the SAME step in one process alternates 159.5 / 162.5 us (tools/placement2.py): is
it the address of the tiny loss buffers capture() re-creates?   python tools/placement3.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from yolo_dl_b200 import synth
from yolo_dl_b200.loss import YoloLossInit, pack_labels
from yolo_dl_b200.pipeline import DeviceStep


def timeit(step, n=300):
    for _ in range(10):
        step.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            step.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / n * 1e3)
    return min(ts)


def main():
    dev = torch.device("cuda:0")
    spec = synth.yolov4_608()
    B = 64
    gt, off = pack_labels(synth.make_targets(spec, B, 20261105))
    gt, off = gt.to(dev), off.to(dev)
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 20261105)]
    step = DeviceStep(spec, B, dev, loss=YoloLossInit(box_metric="ciou"))
    real_zeros = torch.zeros
    print("A: plain re-capture")
    for i in range(6):
        step.capture(heads, gt, off)
        print(f"  {i}: {timeit(step):7.2f} us  losses4@{step.losses4.data_ptr():#x} M@{step.M.data_ptr():#x} K@{step.K.data_ptr():#x}", flush=True)
    print("B: the old loss buffers kept alive (new addresses every time)")
    hold = []
    for i in range(6):
        hold += [step._l5_ring, step._w5_ring]
        step.capture(heads, gt, off)
        print(f"  {i}: {timeit(step):7.2f} us  losses4@{step.losses4.data_ptr():#x}", flush=True)
    print("C: the same loss buffers every time")
    fixed = [real_zeros(5, dtype=torch.float32, device=dev) for _ in range(2)]
    it = iter(())

    def fake_zeros(*a, **k):
        nonlocal it
        if a and a[0] == 5:
            try:
                t = next(it)
                t.zero_()
                return t
            except StopIteration:
                pass
        return real_zeros(*a, **k)
    for i in range(6):
        it = iter(fixed)
        torch.zeros = fake_zeros
        try:
            step.capture(heads, gt, off)
        finally:
            torch.zeros = real_zeros
        print(f"  {i}: {timeit(step):7.2f} us  losses4@{step.losses4.data_ptr():#x}", flush=True)


if __name__ == "__main__":
    main()
