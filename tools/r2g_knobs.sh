#!/bin/bash
# the remaining tunables under the final schedule, one box, one call
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
for v in "YH_X=0" "YH_FILTER_PER=2" "YH_FILTER_PER=4" "YH_DENSE_CTAS_PER_SM=1" "YH_DENSE_CTAS_PER_SM=4" "YH_KEEP_MB=32" "YH_KEEP_MB=96" "YH_DECODE_PER=8" "YH_X=0"; do
  echo -n "$v: "
  env $v timeout 100 python bench.py $Q 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('us/step',round(d['ms_per_step']*1e3,2), [round(x*1e3,2) for x in d['value_windows']['ms_per_step']])
"
done
