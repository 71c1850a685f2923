"""scratch: yh_nms_by_scores with ONE group (per-image NMS), large K"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import yolo_dl_b200 as yh
from yolo_dl_b200 import synth
dev = torch.device("cuda:0")
for K in (4000, 16000, 65536):
    ltrb, scores, grp = synth.make_nms_boxes(K, 7 + K, num_groups=1)
    a = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
    kept = yh.nms_by_scores(*a, 0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        kept = yh.nms_by_scores(*a, 0.5)
    e1.record()
    torch.cuda.synchronize()
    print(f"one group K={K:6d} ms={e0.elapsed_time(e1)/3:.3f} kept={kept.numel()}")
