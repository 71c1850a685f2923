#!/bin/bash
# nms_segment emits by itself (no nms_emit launch): correctness, then A/B against YH_NMS_FUSE_EMIT=0
echo "== correctness"
timeout 600 python -m pytest tests/test_gpu_nms.py tests/test_gpu_pipeline.py tests/test_gpu_golden.py tests/test_gpu_abi_c.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4
timeout 600 python -m pytest tests/test_gpu_aten.py -m gpu -q -x -p no:cacheprovider -k nms 2>&1 | tail -3
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
for B in 64 8; do
for v in "YH_NMS_FUSE_EMIT=0" "YH_X=0" "YH_NMS_FUSE_EMIT=0" "YH_X=0"; do
  echo "== B=$B $v"
  env $v timeout 200 python bench.py --batch $B $Q 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('us/step',round(d['ms_per_step']*1e3,2), [round(x*1e3,2) for x in d['value_windows']['ms_per_step']], 'launches/step', d['gpu_launches']//d['steps']//d['value_windows']['windows'] if 0 else d['gpu_launches'])
"
done; done
echo "== timeline B=64"; timeout 200 python tools/timeline.py --batch 64 --replays 20 2>&1 | tail -11
echo "== timeline B=8"; timeout 200 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -11
