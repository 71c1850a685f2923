#!/usr/bin/env python
"""Is the step's slow mode (165 instead of 158 us, same build, another process) a matter of WHERE
the buffers lie?  One process, several DeviceSteps on fresh allocations (the old ones kept alive
so that the allocator has to hand out other addresses), each captured and timed.

  python tools/placement.py [--rounds 6] [--fresh-heads]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from yolo_dl_b200 import synth
from yolo_dl_b200.loss import YoloLossInit, pack_labels
from yolo_dl_b200.pipeline import DeviceStep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=6)
    ap.add_argument("--fresh-heads", action="store_true")
    ap.add_argument("--replays", type=int, default=400)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    spec = synth.yolov4_608()
    B = 64
    gt, off = pack_labels(synth.make_targets(spec, B, 20261105))
    gt, off = gt.to(dev), off.to(dev)
    keep = []
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 20261105)]
    for r in range(a.rounds):
        if a.fresh_heads and r:
            keep.append(heads)
            heads = [h.clone() for h in heads]
        step = DeviceStep(spec, B, dev, loss=YoloLossInit(box_metric="ciou"))
        step.capture(heads, gt, off)
        for _ in range(10):
            step.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.replays):
                step.replay()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / a.replays * 1e3)
        print(f"round {r}: {min(ts):7.2f} us per step  heads@{heads[0].data_ptr():#x} grads@{step.grads[0].data_ptr():#x} "
              f"boxes@{step.boxes.data_ptr() if hasattr(step.boxes, 'data_ptr') else step.boxes[0].data_ptr():#x}", flush=True)
        keep.append(step)
        # a spacer so that the next round's buffers do not land where a freed block was
        keep.append(torch.empty((37 + 11 * r) << 20, dtype=torch.uint8, device=dev))


if __name__ == "__main__":
    main()
