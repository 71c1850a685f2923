#!/bin/bash
# A/B: end of the dense store stream kept in the L2 (YH_KEEP_MB) + two-sweep scatter, B=64
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
for v in "YH_KEEP_MB=0" "YH_KEEP_MB=32" "YH_KEEP_MB=64" "YH_KEEP_MB=96" "YH_KEEP_MB=120" "YH_KEEP_MB=0" "YH_KEEP_MB=96"; do
  echo "== $v"
  env $v timeout 200 python bench.py $Q 2>gpurun_out/keep.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],5), [round(x,5) for x in d['value_windows']['ms_per_step']], 'dense_ms',round(r['kernel_ms'],4),'step_frac',round(r['step']['frac'],3))
print({k:round(v['ms'],4) for k,v in r['other_kernels'].items()})
print('in_graph', r.get('in_graph', {}).get('span_us'), {k: v['us_sum'] for k, v in r.get('in_graph', {}).get('per_kernel', {}).items()})
"; tail -2 gpurun_out/keep.err
done
echo "== correctness with YH_KEEP_MB=96"
YH_KEEP_MB=96 timeout 900 python -m pytest tests/test_gpu_loss.py tests/test_gpu_pipeline.py tests/test_gpu_aten.py tests/test_gpu_golden.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -5
