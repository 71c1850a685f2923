#!/bin/bash
# round-2 GPU session 1: large-segment NMS timing, the whole GPU suite, schedule A/B
mkdir -p gpurun_out
timeout 300 python tools/nms_single_group.py 1 > gpurun_out/nms1.log 2>&1; echo "nms1 rc=$?"; tail -8 gpurun_out/nms1.log
timeout 2400 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_gpu.log
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-e2e --no-extras --no-strong --windows 3"
for cfg in "YH_SCHEDULE=overlapped" "YH_SCHEDULE=grouped YH_GROUP=8 YH_DENSE_AFTER=assign" "YH_SCHEDULE=grouped YH_GROUP=8 YH_DENSE_AFTER=fork" "YH_SCHEDULE=grouped YH_GROUP=16 YH_DENSE_AFTER=assign" "YH_SCHEDULE=grouped YH_GROUP=4 YH_DENSE_AFTER=assign"; do
  echo "== $cfg"
  env $cfg timeout 300 $B 2> gpurun_out/bench_err.log | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'step_frac', d['roofline']['step']['frac'], 'launches', d['gpu_launches'], 'loss', d['loss'])
    print({k: round(v['ms'], 4) for k, v in d['roofline']['other_kernels'].items()}, d['roofline']['kernel_ms'])
"
  tail -3 gpurun_out/bench_err.log
done
