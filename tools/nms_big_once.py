"""one yh_nms_by_scores call on ONE group of 65 536 rows (ncu target for nms_big_kernel)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import yolo_dl_b200 as yh
from yolo_dl_b200 import synth
dev = torch.device("cuda:0")
ltrb, scores, grp = synth.make_nms_boxes(65536, 7 + 65536, num_groups=1)
a = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
for _ in range(2):
    kept = yh.nms_by_scores(*a, 0.5)
torch.cuda.synchronize()
print("kept", kept.numel())
