#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_loss.py tests/test_gpu_nms.py tests/test_gpu_golden.py tests/test_gpu_benchmark.py -m gpu -q -p no:cacheprovider 2>&1 | tail -4 | cut -c1-200
timeout 200 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -12
python - <<'PY'
import sys, os, statistics, json
sys.path.insert(0, os.getcwd())
import torch, bench
from yolo_dl_b200 import synth
dev = torch.device("cuda:0")
spec = synth.yolov4_608()
for b in (8, 16, 32, 64):
    us = [bench.step_us_at_batch(dev, spec, b, 300, 20261200 + b)[0] for _ in range(2)]
    print("batch", b, "us_per_step", [round(x, 2) for x in us])
PY
