#!/bin/bash
mkdir -p gpurun_out
echo skip tests
B="python bench.py --steps 400 --warmup 10 --no-cpu-baseline --no-e2e --no-extras --no-strong --windows 3"
for rep in 1 2; do
for cfg in "YH_FINISH_EARLY=0" "YH_FINISH_EARLY=1" "YH_FINISH_EARLY=1 YH_FINISH_EARLY_UNCAPPED=1"; do
  echo "== rep $rep $cfg"
  env $cfg timeout 300 $B 2> gpurun_out/bench_err.log | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('ms_per_step', round(d['ms_per_step'],5), 'step_frac', round(d['roofline']['step']['frac'],4), 'loss', d['loss'])
    ig = d['roofline'].get('in_graph', {})
    print('in_graph', ig.get('span_us'), [(l['kernel'][:12], l['start_us'], l['end_us']) for l in ig.get('launches', []) if 'finish' in l['kernel'] or 'dense' in l['kernel'] or 'rows' in l['kernel']])
"
  tail -2 gpurun_out/bench_err.log | cut -c1-200
done
done
