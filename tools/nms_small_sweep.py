"""scratch: yh_nms_by_scores on small inputs, single-segment vs bucketed path (YH_BY_MIN)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import yolo_dl_b200 as yh
from yolo_dl_b200 import synth
dev = torch.device("cuda:0")
for groups in (80, 1):
    for K in (128, 256, 512, 1000, 2000, 4000):
        ltrb, scores, grp = synth.make_nms_boxes(K, 7 + K, num_groups=groups)
        a = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
        kept = yh.nms_by_scores(*a, 0.5)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            kept = yh.nms_by_scores(*a, 0.5)
        e1.record()
        torch.cuda.synchronize()
        print(f"BY_MIN={os.environ.get('YH_BY_MIN','2048'):>5} groups={groups:3d} K={K:5d} ms={e0.elapsed_time(e1)/20:.4f} kept={kept.numel()}")
