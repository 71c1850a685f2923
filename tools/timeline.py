#!/usr/bin/env python
"""Where does each kernel run inside the overlapped, graph-replayed step?

  python tools/timeline.py [--batch 64] [--replays 20]

Captures DeviceStep's graph with yh_trace_enable on (a %globaltimer stamp before and after every
launch, inside the graph), replays it and prints start/end of every kernel relative to the
earliest start, median over the replays.  The stamps cost two extra one-thread nodes per kernel,
so the absolute step time is a few microseconds longer than bench.py's.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from yolo_dl_b200 import _lib, synth
from yolo_dl_b200.loss import YoloLossInit, pack_labels
from yolo_dl_b200.pipeline import DeviceStep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--replays", type=int, default=20)
    ap.add_argument("--sequential", action="store_true", help="one stream instead of the fork/join")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    spec = synth.yolov4_608()
    B = a.batch
    heads = [h.to(dev) for h in synth.make_heads(spec, B, 20261105)]
    targets = synth.make_targets(spec, B, 20261105)
    gt, off = pack_labels(targets)
    gt, off = gt.to(dev), off.to(dev)
    step = DeviceStep(spec, B, dev, loss=YoloLossInit(box_metric="ciou"))
    import ctypes
    nk = 64
    buf = torch.zeros(2 * nk, dtype=torch.int64, device=dev)
    if not a.sequential:
        step.capture(heads, gt, off)      # warm-up capture without tracing
    lib.yh_trace_enable(buf.data_ptr(), nk)
    if a.sequential:
        step.run(heads, gt, off)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        cap = torch.cuda.Stream(dev)
        with torch.cuda.graph(g, stream=cap):
            step.run(heads, gt, off, stream=cap.cuda_stream)
        replay = g.replay
    else:
        step.capture(heads, gt, off)
        replay = step.replay
    ids = (ctypes.c_int32 * nk)()
    n = min(nk, lib.yh_trace_ids(ids, nk))
    lib.yh_trace_enable(None, 0)
    # the warm-up run inside capture() was traced too: the graph's launches are the last ones
    k = n // 2 if a.sequential else step.kernels_in_graph
    first = n - k
    names = [_lib.KERNEL_NAMES[ids[i]] for i in range(first, n)]
    rows = []
    for _ in range(a.replays):
        replay()
        torch.cuda.synchronize()
        rows.append(buf.cpu().numpy().astype(np.int64).reshape(nk, 2))
    t = np.stack(rows[3:])[:, first:n]  # [R, k, 2]
    n = k
    t0 = t[:, :, 0].min(axis=1)[:, None, None]
    rel = np.median((t - t0) / 1000.0, axis=0)
    order = sorted(range(n), key=lambda i: rel[i, 0])
    print(f"{'kernel':24s} {'start us':>9s} {'end us':>9s} {'dur us':>8s}")
    for i in order:
        print(f"{names[i]:24s} {rel[i,0]:9.1f} {rel[i,1]:9.1f} {rel[i,1]-rel[i,0]:8.1f}")
    print("span us:", float(rel[:, 1].max()))


if __name__ == "__main__":
    main()
