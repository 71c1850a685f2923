# scratch A/B driver: bash tools/ab.sh "<env assignments>" ...   (one bench.py run per argument)
mkdir -p gpurun_out; rm -f gpurun_out/ab.txt gpurun_out/ab.err
Q="--steps 400 --warmup 10 --no-e2e --no-cpu-baseline --no-extras"
for v in "$@"; do
  echo "== $v" >> gpurun_out/ab.txt
  env $v python bench.py $Q 2>>gpurun_out/ab.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('value',round(d['value']),'ms/step',round(d['ms_per_step'],4),'seq',round(d['sequential_ms_per_step'],4),'dense_ms',round(r['kernel_ms'],4),'frac',round(r['frac'],3))
print({k:round(v['ms'],4) for k,v in r['other_kernels'].items()})
" >> gpurun_out/ab.txt 2>&1
done
cat gpurun_out/ab.txt
