#!/bin/bash
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
for v in "YH_X=0" "YH_DEC_PRIO=-1" "YH_NMS_PRIO=0" "YH_NMS_PRIO=0 YH_DEC_PRIO=-1" "YH_X=0"; do
  echo "== $v"
  env $v timeout 200 python bench.py $Q 2>gpurun_out/prio.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],5), [round(x,5) for x in d['value_windows']['ms_per_step']], 'step_frac',round(r['step']['frac'],3))
print('in_graph', r.get('in_graph', {}).get('span_us'), {k: v['us_sum'] for k, v in r.get('in_graph', {}).get('per_kernel', {}).items()})
"; tail -2 gpurun_out/prio.err
done
