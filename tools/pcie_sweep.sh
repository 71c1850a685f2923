#!/bin/bash
# tools/pcie_sweep.sh <max_gpus>: bare pinned H2D || D2H of 495 MB per GPU from 1/2/4/8 concurrent
# processes, unbound and NUMA-bound (is 8.2 GB/s per GPU at N=8 the box or the
# pipeline?).  Output: gpurun_out/pcie.txt
N=${1:-8}
OUT=gpurun_out/pcie.txt
mkdir -p gpurun_out
{
  echo "# $(date -u) host: $(nproc) cpus; numa nodes: $(ls -d /sys/devices/system/node/node* 2>/dev/null | wc -l)"
  nvidia-smi topo -m 2>/dev/null | head -20
  lscpu 2>/dev/null | grep -i "model name\|socket\|numa"
  for bind in 0 1; do
    for n in 1 2 4 8; do
      [ $n -gt $N ] && continue
      echo "== $n process(es), bind=$bind"
      start=$(python3 -c "import time; print(time.time() + 6)")
      for d in $(seq 0 $((n-1))); do
        tools/pcie_bench $d 495 10 $start $bind &
      done
      wait
    done
  done
} > $OUT 2>&1
cat $OUT
