#!/bin/bash
# small batches, same box: nothing kept (the new rule) against the whole gradient kept + two sweeps
for B in 8 16 32; do
for v in "YH_X=0" "YH_KEEP_MB=1" "YH_X=0" "YH_KEEP_MB=1"; do
  echo "== B=$B $v"
  env $v timeout 200 python bench.py --batch $B --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('us/step',round(d['ms_per_step']*1e3,2), [round(x*1e3,2) for x in d['value_windows']['ms_per_step']])
"
done; done
