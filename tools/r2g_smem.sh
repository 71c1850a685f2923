#!/bin/bash
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
run() {
  echo "== $1"
  env $1 timeout 200 python bench.py $Q 2>gpurun_out/smem.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],5), [round(x,5) for x in d['value_windows']['ms_per_step']], 'step_frac',round(r['step']['frac'],3), 'dense alone', round(r['kernel_ms'],4))
print('in_graph', r.get('in_graph', {}).get('span_us'), {k: v['us_sum'] for k, v in r.get('in_graph', {}).get('per_kernel', {}).items()})
"; tail -2 gpurun_out/smem.err
}
run "YH_X=0"
for kb in 28 32 38 46; do
  run "YH_DENSE_SMEM_KB=$kb"
  run "YH_DENSE_SMEM_KB=$kb YH_DENSE_LAST=0"
  run "YH_DENSE_SMEM_KB=$kb YH_DENSE_LAST=0 YH_NMS_PRIO=-1 YH_DEC_PRIO=0"
  run "YH_DENSE_SMEM_KB=$kb YH_DENSE_LAST=0 YH_NMS_PRIO=0 YH_DEC_PRIO=0 YH_ASG_PRIO=0"
done
echo "== timeline 46 dense first"
YH_DENSE_SMEM_KB=46 YH_DENSE_LAST=0 timeout 200 python tools/timeline.py --batch 64 --replays 20 2>&1 | tail -12
