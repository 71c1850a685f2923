#!/bin/bash
# loss_finish_kernel with a register cap (min blocks per SM 4 / 6 / 8) against the shipped build
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras --no-strong --windows 3"
for lib in libyolohot.so libyolohot_f4.so libyolohot_f6.so libyolohot_f8.so libyolohot.so libyolohot_f4.so; do
  echo "== $lib"
  YH_LIB=$PWD/yolo_dl_b200/$lib timeout 200 python bench.py $Q 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('us/step',round(d['ms_per_step']*1e3,2), [round(x*1e3,2) for x in d['value_windows']['ms_per_step']], 'finish alone', round(d['roofline']['other_kernels']['loss_finish_kernel']['ms']*1e3,1), 'loss', d['loss'])
"
done
