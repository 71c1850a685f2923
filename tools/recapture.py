#!/usr/bin/env python
"""The same step captured N times in one process: spread of the step time over the captures.
  [env ...] python tools/recapture.py [N]"""
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
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    dev = torch.device("cuda:0")
    spec = synth.yolov4_608()
    B = int(os.environ.get("B", "64"))
    gt, off = pack_labels(synth.make_targets(spec, B, 20261105))
    gt, off = gt.to(dev), off.to(dev)
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 20261105)]
    step = DeviceStep(spec, B, dev, loss=YoloLossInit(box_metric="ciou"))
    out = []
    for i in range(n):
        step.capture(heads, gt, off)
        out.append(timeit(step))
    print(" ".join(f"{t:.1f}" for t in out), "| min %.1f max %.1f" % (min(out), max(out)), flush=True)


if __name__ == "__main__":
    main()
