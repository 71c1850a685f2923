"""scratch: per-kernel times of yh_nms_by_scores at large K (80 groups)"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import yolo_dl_b200 as yh
from yolo_dl_b200 import synth, _lib
dev = torch.device("cuda:0")
lib = yh.load()
for K in (65536, 200000):
    ltrb, scores, grp = synth.make_nms_boxes(K, 7 + K, num_groups=80)
    a = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
    for _ in range(2):
        kept = yh.nms_by_scores(*a, 0.5)
    torch.cuda.synchronize()
    lib.yh_timing_enable(1)
    for _ in range(5):
        kept = yh.nms_by_scores(*a, 0.5)
    torch.cuda.synchronize()
    nk = len(_lib.KERNEL_NAMES)
    sums = (ctypes.c_double * nk)(); cnts = (ctypes.c_int64 * nk)()
    _lib.check(lib.yh_timing_read(sums, cnts, nk))
    lib.yh_timing_enable(0)
    print(K, {n: (round(sums[i] / 5, 3), cnts[i] // 5) for i, n in enumerate(_lib.KERNEL_NAMES) if cnts[i]})
