"""scratch: yh_nms_by_scores with ONE group (per-image NMS), per-kernel times"""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import yolo_dl_b200 as yh
from yolo_dl_b200 import synth, _lib
dev = torch.device("cuda:0")
lib = yh.load()
groups = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for K in (1000, 4000, 5000, 16000, 65536, 100000, 200000):
    ltrb, scores, grp = synth.make_nms_boxes(K, 7 + K, num_groups=groups)
    a = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
    for _ in range(2):
        kept = yh.nms_by_scores(*a, 0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        kept = yh.nms_by_scores(*a, 0.5)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    lib.yh_timing_enable(1)
    for _ in range(3):
        kept = yh.nms_by_scores(*a, 0.5)
    torch.cuda.synchronize()
    nk = len(_lib.KERNEL_NAMES)
    sums = (ctypes.c_double * nk)(); cnts = (ctypes.c_int64 * nk)()
    _lib.check(lib.yh_timing_read(sums, cnts, nk))
    lib.yh_timing_enable(0)
    print(f"groups={groups} K={K} kept={kept.numel()} {ms:.3f} ms/call",
          {n: (round(sums[i] / 3, 3), cnts[i] // 3) for i, n in enumerate(_lib.KERNEL_NAMES) if cnts[i]}, flush=True)
