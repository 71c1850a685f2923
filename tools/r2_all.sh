#!/bin/bash
# one acquisition, everything that needs one GPU and no profiler: NMS timing, the whole GPU suite,
# schedule A/B, timelines, full bench line.  Every stage under its own timeout; logs in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "### 1. single-group NMS"; timeout 300 python tools/nms_single_group.py 1 > gpurun_out/nms1.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/nms1.log; timeout 300 python tools/nms_single_group.py 80 > gpurun_out/nms80.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/nms80.log
echo "### 2. smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/smoke.log
echo "### 3. pytest -m gpu"; timeout 2700 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_gpu.log | cut -c1-300
echo "### 4. schedule A/B (B=64)"
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-e2e --no-extras --no-strong --windows 3"
for cfg in "YH_SCHEDULE=overlapped" "YH_SCHEDULE=overlapped" "YH_L2_HINT=0"; do
  echo "== $cfg"
  env $cfg timeout 300 $B 2> gpurun_out/bench_err.log | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('ms_per_step', d['ms_per_step'], d['value_windows']['ms_per_step'], 'step_frac', d['roofline']['step']['frac'], 'launches', d['gpu_launches'], 'loss', d['loss'])
    print({k: round(v['ms'], 4) for k, v in d['roofline']['other_kernels'].items()}, d['roofline']['kernel_ms'])
    print('in_graph', d['roofline'].get('in_graph', {}).get('span_us'), d['roofline'].get('in_graph', {}).get('per_kernel'))
"
  tail -3 gpurun_out/bench_err.log
done
echo "### 5. timelines"
for cfg in "YH_PDL=0" "YH_PDL=1"; do
  echo "== B=8 $cfg"
  env $cfg timeout 200 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -22
done
echo "== B=64 overlapped"; timeout 200 python tools/timeline.py --batch 64 --replays 20 2>&1 | tail -16
echo "### 6. full default bench"
timeout 1200 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_full.err
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches')}, d['value_windows'], d['roofline']['step'])
    print('e2e', d['e2e']); print('strong', d['strong']); print('cpu', d['cpu_baseline'])
    for k, v in (d.get('other_configs') or {}).items(): print(k, v)
except Exception as e:
    print('parse failed', e)
PY
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench_full.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-400
