#!/bin/bash
# B=8 (strong-scaling share of one GPU): overlapped vs "grouped" with one range (rows store straight
# over the dense gradient, no scatter kernel)
Q="--batch 8 --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
for v in "YH_SCHEDULE=overlapped" "YH_SCHEDULE=grouped YH_GROUP=8 YH_DENSE_AFTER=none" "YH_SCHEDULE=grouped YH_GROUP=8 YH_DENSE_AFTER=assign"; do
  echo "== $v"
  env $v timeout 200 python bench.py $Q 2>gpurun_out/b8.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms/step',round(d['ms_per_step'],5), d['value_windows']['ms_per_step'], 'launches', d['gpu_launches'])
" ; tail -2 gpurun_out/b8.err
  env $v timeout 100 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -14
done
