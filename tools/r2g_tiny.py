#!/usr/bin/env python
"""decode + NMS of YOLOv4-tiny-416 at 256 images and of YOLOv4-608 at 8 (BASELINE configs[2], [0]) as
a replayed graph: us per batch.   [YH_NMS_FUSE_EMIT=0] python tools/r2g_tiny.py"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import yolo_dl_b200 as yh
from yolo_dl_b200 import synth
from yolo_dl_b200.pipeline import DeviceStep

dev = torch.device("cuda:0")
for name, spec, B, conf, iou in (("tiny-416 x256", synth.yolov4_tiny_416(), 256, 0.4, 0.5),
                                 ("608 x8", synth.yolov4_608(), 8, 0.4, 0.5), ("608 x64", synth.yolov4_608(), 64, 0.4, 0.5)):
    heads = synth.make_heads(spec, B, 20261107, device=dev)
    hp = yh._lib.ptr_array(heads)
    for by_class in (False, True):
        st = DeviceStep(spec, B, dev, nms_conf=conf, nms_iou=iou, by_class=by_class, with_grad=False)

        def f(s=None, st=st):
            s = s if s is not None else torch.cuda.current_stream(dev).cuda_stream
            st._decode(hp, s)
            st._nms(hp, s)
        us = [bench.graph_us(dev, f) for _ in range(3)]
        print(name, "per_class" if by_class else "per_image", " ".join(f"{u:.1f}" for u in us), "us")
