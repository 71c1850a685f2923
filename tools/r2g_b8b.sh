#!/bin/bash
Q="--batch 8 --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
echo "== correctness: NMS + pipeline tests"
timeout 900 python -m pytest tests/test_gpu_nms.py tests/test_gpu_pipeline.py tests/test_gpu_golden.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4
timeout 900 python -m pytest tests/test_gpu_aten.py -m gpu -q -x -p no:cacheprovider -k nms 2>&1 | tail -3
YH_ROWS_ON_MAIN=1 timeout 900 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
for v in "YH_X=0" "YH_ROWS_ON_MAIN=1" "YH_X=0" "YH_ROWS_ON_MAIN=1 YH_PDL=0"; do
  echo "== $v"
  env $v timeout 200 python bench.py $Q 2>gpurun_out/b8.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms/step',round(d['ms_per_step'],5), d['value_windows']['ms_per_step'], 'launches', d['gpu_launches'])
" ; tail -2 gpurun_out/b8.err
done
for v in "YH_X=0" "YH_ROWS_ON_MAIN=1"; do
  echo "== timeline B=8 $v"
  env $v timeout 100 python tools/timeline.py --batch 8 --replays 30 2>&1 | tail -12
done
echo "== segprof"
bash tools/r2g_segprof.sh
echo "== B=64 check"
timeout 200 python bench.py --steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms/step',round(d['ms_per_step'],5), d['value_windows']['ms_per_step'])"
