#!/bin/bash
# phase stamps of nms_segment_kernel (profiling build) inside the B=8 and B=64 step
for B in 8 64; do
  echo "== B=$B"
  YH_LIB=$PWD/yolo_dl_b200/libyolohot_prof.so timeout 200 python tools/timeline.py --batch $B --replays 8 2>&1 | grep SEGPROF | tail -6
done
