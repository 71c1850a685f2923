"""Scratch probe: per-kernel CUDA-event times of the step for a few settings (run on the GPU box)."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import yolo_dl_b200 as yh
from yolo_dl_b200 import _lib, synth
from yolo_dl_b200.pipeline import DeviceStep

dev = torch.device("cuda:0")
lib = yh.load()
spec = synth.yolov4_608()


def probe(B, conf, by_class, n_max=30, reps=20):
    heads = synth.make_heads(spec, B, 1, device=dev)
    targets = synth.make_targets(spec, B, 1, n_max=n_max)
    gt, off = yh.pack_labels(targets)
    gt, off = gt.to(dev), off.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=conf, nms_iou=0.5, by_class=by_class, cand_per_image=65536, max_detections=B * 65536)
    for _ in range(3):
        step.run(heads, gt, off)
    M, K = step.counts()
    cand = int(torch.as_tensor(0))
    lib.yh_timing_enable(1)
    for _ in range(reps):
        step.run(heads, gt, off)
    sums = (ctypes.c_double * 10)(); cnts = (ctypes.c_int64 * 10)()
    lib.yh_timing_read(sums, cnts, 10)
    lib.yh_timing_enable(0)
    t = {n[:-7]: round(sums[i] / cnts[i] * 1e3, 1) for i, n in enumerate(_lib.KERNEL_NAMES) if cnts[i]}
    print(f"B={B} conf={conf} by_class={by_class} n_max={n_max} M={M} K={K} us:", t, flush=True)


for B in (1, 8, 64):
    probe(B, 0.4, True)
probe(64, 0.99, True)
probe(64, 0.1, True)
probe(64, 0.4, False)
probe(64, 0.4, True, n_max=2)
