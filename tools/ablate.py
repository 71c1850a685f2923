#!/usr/bin/env python
"""What does each branch of the overlapped step COST the step?  (diagnosis, not a benchmark)

  python tools/ablate.py [--batch 64] [--replays 300]

Captures DeviceStep's graph with one branch left out at a time (the C-ABI call replaced by a
no-op) and times the replays: the difference to the full step is that branch's marginal cost —
the kernels' alone-times say little, because the memory-bound ones do not overlap each other.
The ablated steps compute wrong results by construction; nothing here is a reported number.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from yolo_dl_b200 import _lib, synth
from yolo_dl_b200.loss import YoloLossInit, pack_labels
from yolo_dl_b200.pipeline import DeviceStep


class _Without:
    """the library with some entry points turned into no-ops"""

    def __init__(self, lib, names):
        self._lib, self._names = lib, set(names)

    def __getattr__(self, k):
        if k in self._names:
            return lambda *a: 0
        return getattr(self._lib, k)


CASES = {
    "full": [],
    "-decode": ["yh_decode_merged"],
    "-nms": ["yh_nms"],
    "-decode -nms": ["yh_decode_merged", "yh_nms"],
    "-scatter (finish)": ["yh_loss_finish"],
    "-rows -scatter": ["yh_loss_rows_compute", "yh_loss_finish"],
    "-assign -rows -scatter": ["yh_assign", "yh_loss_rows_compute", "yh_loss_finish"],
    "-dense": ["yh_loss_dense"],
    "dense only": ["yh_decode_merged", "yh_nms", "yh_assign", "yh_loss_rows_compute", "yh_loss_finish"],
    "dense + scatter": ["yh_decode_merged", "yh_nms", "yh_assign", "yh_loss_rows_compute"],
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--replays", type=int, default=300)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    spec = synth.yolov4_608()
    B = a.batch
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 20261105)]
    gt, off = pack_labels(synth.make_targets(spec, B, 20261105))
    gt, off = gt.to(dev), off.to(dev)
    real = _lib.load()
    for name, drop in CASES.items():
        step = DeviceStep(spec, B, dev, loss=YoloLossInit(box_metric="ciou"))
        # one full step first, so that the match list the scatter reads exists in every case
        step.run_overlapped(heads, gt, off)
        torch.cuda.synchronize()
        step.lib = _Without(real, drop)
        step.capture(heads, gt, off)
        for _ in range(10):
            step.replay()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.replays):
                step.replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / a.replays * 1e3)
        print(f"{name:28s} {best:8.1f} us per step  ({step.kernels_in_graph} launches)")
        del step


if __name__ == "__main__":
    main()
