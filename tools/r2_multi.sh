#!/bin/bash
# tools/r2_multi.sh <N>: what needs more than one GPU.  N=2: the two-thread / two-device test, the
# NCCL gradient all-reduce test, the peer-mailbox test.  N=8: bare PCIe sweep, then bench.py at 2/4/8.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
if [ "$N" = 2 ]; then
  timeout 900 python -m pytest tests/test_gpu_multidev.py tests/test_gpu_peer.py -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_2gpu.log 2>&1
  echo "2-gpu pytest rc=$?"; tail -6 gpurun_out/pytest_2gpu.log | cut -c1-200
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 200 --warmup 5 --windows 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
  echo "bench n2 rc=$?"; tail -2 gpurun_out/bench_n2.err | cut -c1-200
else
  make -C yolo_dl_b200/csrc pcie_bench > /dev/null 2>&1
  bash tools/pcie_sweep.sh 8 > /dev/null 2>&1; tail -40 gpurun_out/pcie.txt | cut -c1-200
  for n in 8 4 2; do
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 200 --warmup 5 --windows 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
    echo "bench n$n rc=$?"; tail -2 gpurun_out/bench_n$n.err | cut -c1-200
  done
fi
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/bench_n*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value']), 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None,
              'pcie', d['e2e'].get('pcie_GBps_each_way') if d.get('e2e') else None)
        print('  strong', {k: d['strong'][k] for k in ('batch_per_gpu', 'value', 'ms_per_step')} if d.get('strong') else None)
        print('  dp', d.get('dp_gradient_allreduce'))
        print('  exchange', d.get('loss_exchange'))
    except Exception as e:
        print(f, 'parse failed', e)
PY
