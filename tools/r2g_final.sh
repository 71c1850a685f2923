#!/bin/bash
# r2g: one acquisition, no profiler — smoke, the whole GPU suite, old-vs-new schedule A/B on the
# same box, timelines, the full default bench line and the reference arm.  Logs in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
echo "### 1. smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2g_smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r2g_smoke.log
echo "### 2. pytest -m gpu"; timeout 2700 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2g_pytest_gpu.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2g_pytest_gpu.log | cut -c1-300
echo "### 3. schedule A/B (B=64): r2f settings, then the new defaults, twice"
B="python bench.py --steps 400 --warmup 10 --no-cpu-baseline --no-e2e --no-extras --no-strong --windows 3"
for cfg in "YH_DEC_PRIO=0 YH_NMS_PRIO=-1 YH_KEEP_MB=0 YH_NMS_FUSE_EMIT=0 YH_DENSE_AFTER_DECODE=0" "YH_X=0" "YH_DEC_PRIO=0 YH_NMS_PRIO=-1 YH_KEEP_MB=0 YH_NMS_FUSE_EMIT=0 YH_DENSE_AFTER_DECODE=0" "YH_X=0"; do
  echo "== $cfg"
  env $cfg timeout 300 $B 2> gpurun_out/bench_err.log | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('ms_per_step', d['ms_per_step'], d['value_windows']['ms_per_step'], 'step_frac', d['roofline']['step']['frac'], 'clocks', d.get('clocks'))
    print({k: round(v['ms'], 4) for k, v in d['roofline']['other_kernels'].items()}, d['roofline']['kernel_ms'])
"
  tail -3 gpurun_out/bench_err.log
done
echo "### 4. timelines"
echo "== B=8"; timeout 200 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -12
echo "== B=64"; timeout 200 python tools/timeline.py --batch 64 --replays 20 2>&1 | tail -12
echo "### 5. full default bench"
timeout 1500 python bench.py > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2g_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/r2g_bench.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches')}, d['value_windows'], d['roofline']['step'])
    print('e2e', d['e2e']); print('strong', d['strong']); print('cpu', d['cpu_baseline'])
    for k, v in (d.get('other_configs') or {}).items(): print(k, v)
except Exception as e:
    print('parse failed', e)
PY
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2g_bench_reference.json 2>> gpurun_out/r2g_bench.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/r2g_bench_reference.json
