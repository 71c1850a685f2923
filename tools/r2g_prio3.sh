#!/bin/bash
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
run() {
  echo "== $1"
  env $1 timeout 200 python bench.py $Q 2>gpurun_out/prio.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],5), [round(x,5) for x in d['value_windows']['ms_per_step']], 'step_frac',round(r['step']['frac'],3))
print('in_graph', r.get('in_graph', {}).get('span_us'), {k: v['us_sum'] for k, v in r.get('in_graph', {}).get('per_kernel', {}).items()})
"; tail -2 gpurun_out/prio.err
}
run "YH_NMS_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2"
run "YH_NMS_PRIO=-1 YH_FILTER_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2"
run "YH_NMS_PRIO=-2 YH_FILTER_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2"
run "YH_NMS_PRIO=-1 YH_FILTER_PRIO=0 YH_DEC_PRIO=-1 YH_ASG_PRIO=-1"
run "YH_NMS_PRIO=-1 YH_FILTER_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2 YH_FILTER_PER=2"
run "YH_NMS_PRIO=-1 YH_FILTER_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2 YH_FILTER_PER=4"
run "YH_NMS_PRIO=-1 YH_FILTER_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2 YH_KEEP_MB=0"
echo "== timeline"
YH_NMS_PRIO=-1 YH_FILTER_PRIO=0 YH_DEC_PRIO=-3 YH_ASG_PRIO=-2 timeout 200 python tools/timeline.py --batch 64 --replays 20 2>&1 | tail -12
echo "== timeline B=8 filter prio 0, others default"
YH_FILTER_PRIO=0 timeout 200 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -12
