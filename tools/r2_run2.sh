#!/bin/bash
# round-2 GPU session 2: where the time goes at 8 images per GPU (timelines, PDL on/off), the L2
# prefetch-size hint, and the full default bench line
mkdir -p gpurun_out
for cfg in "YH_SCHEDULE=overlapped YH_PDL=0" "YH_SCHEDULE=overlapped YH_PDL=1" "YH_SCHEDULE=grouped YH_DENSE_AFTER=fork YH_PDL=1"; do
  echo "== B=8 $cfg"
  env $cfg timeout 200 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -22
done
echo "== B=64 grouped timeline"
timeout 200 python tools/timeline.py --batch 64 --replays 20 2>&1 | tail -34
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-e2e --no-extras --no-strong --windows 3"
for cfg in "YH_L2_HINT=0" "YH_L2_HINT=1"; do
  echo "== $cfg"
  env $cfg timeout 300 $B 2> gpurun_out/bench_err.log | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('ms_per_step', d['ms_per_step'], 'step_frac', d['roofline']['step']['frac'])
    print({k: round(v['ms'], 4) for k, v in d['roofline']['other_kernels'].items()}, d['roofline']['kernel_ms'])
"
  tail -3 gpurun_out/bench_err.log
done
echo "== full default bench"
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches')}, d['value_windows'], d['roofline']['step'])
    print('e2e', d['e2e']); print('strong', d['strong']); print('cpu', d['cpu_baseline'])
    print('in_graph', d['roofline'].get('in_graph', {}).get('per_kernel'))
    for k, v in (d.get('other_configs') or {}).items(): print(k, v)
except Exception as e:
    print('parse failed', e)
PY
