#!/bin/bash
# the round's remaining GPU calls, one after another (each its own gpurun call with retries)
cd "$(dirname "$0")/.."
tools/gpu_retry.sh 2700 'bash tools/r2_sanitize.sh memcheck' > gpurun_out/chain_memcheck.txt 2>&1
tools/gpu_retry.sh 2700 'bash tools/r2_sanitize.sh racecheck' > gpurun_out/chain_racecheck.txt 2>&1
tools/gpu_retry.sh 2700 'bash tools/r2_sanitize.sh synccheck' > gpurun_out/chain_synccheck.txt 2>&1
tools/gpu_retry.sh 1800 'bash tools/r2_ncu.sh' > gpurun_out/chain_ncu.txt 2>&1
/usr/local/graft/bin/gpurun --gpus 2 --timeout 1500 -- 'bash tools/r2_multi.sh 2' > gpurun_out/chain_2gpu.txt 2>&1
