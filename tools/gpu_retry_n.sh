#!/bin/bash
# usage: tools/gpu_retry_n.sh <gpus> <timeout_s> <command...>
N=$1; T=$2; shift; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus $N --timeout $T -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
exit 3
