#!/bin/bash
# builds variants of libyolohot.so that differ in nms.cu compile-time switches (run HERE, no GPU):
#   tools/ab_libs.sh build      -> yolo_dl_b200/libyolohot_<tag>.so for every tag below
# and on the GPU box:  tools/ab_libs.sh run   -> bench.py once per library, same box, same session
cd "$(dirname "$0")/.."
TAGS="oldwalk:-DYH_SEG_OLD_WALK norank:-DYH_SEG_NO_RANKSORT oldboth:-DYH_SEG_OLD_WALK,-DYH_SEG_NO_RANKSORT"
if [ "$1" = build ]; then
  make -C yolo_dl_b200/csrc -j8 all > /dev/null
  cd yolo_dl_b200/csrc
  for t in $TAGS; do
    tag=${t%%:*}; flags=$(echo ${t#*:} | tr ',' ' ')
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 -Xcompiler -fPIC,-Wall -I../../include $flags -c nms.cu -o build/nms_$tag.o
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libyolohot_$tag.so build/ffi.o build/decode.o build/assign.o build/loss.o build/nms_$tag.o build/benchmark.o build/eval.o build/peer.o -lcudart_static -lpthread -ldl -lrt
    echo built $tag
  done
  exit 0
fi
Q="--steps 400 --warmup 10 --windows 3 --no-e2e --no-cpu-baseline --no-extras --no-strong"
for rep in 1 2; do
for lib in "" oldwalk norank oldboth; do
  L=""; [ -n "$lib" ] && L="YH_LIB=$PWD/yolo_dl_b200/libyolohot_$lib.so"
  echo "== rep $rep lib=${lib:-current} $EXTRA"
  env $L $EXTRA timeout 300 python bench.py $Q 2>>gpurun_out/ab.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],4),[round(x,4) for x in d['value_windows']['ms_per_step']],'seq',round(d['sequential_ms_per_step'],4))
print({k:round(v['ms'],4) for k,v in r['other_kernels'].items()})
print({k:v['us_sum'] for k,v in r['in_graph']['per_kernel'].items()}, r['in_graph']['span_us'])
"
done
done
