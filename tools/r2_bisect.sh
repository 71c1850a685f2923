#!/bin/bash
mkdir -p gpurun_out
for cfg in "X=1" "YH_PDL=0" "YH_ROWS_PER_WARP=8" "YH_PDL=0 YH_ROWS_PER_WARP=8"; do
  echo "== $cfg"
  env $cfg timeout 600 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -p no:cacheprovider -k "refilled or host_pipeline_overlapped or host_step_round" 2>&1 | tail -4 | cut -c1-200
done
