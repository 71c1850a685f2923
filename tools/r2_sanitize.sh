#!/bin/bash
# compute-sanitizer over the small-input GPU tests: memcheck on the whole
# "small or golden" selection, racecheck + synccheck on the kernels with hand-rolled shared-memory
# protocols (loss_finish merged stores, nms_segment/nms_big aliased shared memory, assign_emit's
# EMPTY-slot invariant).  Summaries -> gpurun_out/sanitizer_*.txt (copied to profiles/ by hand).
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
SEL='small or golden or abi_smoke'
timeout 1500 $CS --tool memcheck --error-exitcode 9 --print-limit 20 \
  python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "$SEL" > gpurun_out/sanitizer_memcheck.txt 2>&1
echo "memcheck rc=$?"; tail -5 gpurun_out/sanitizer_memcheck.txt
RSEL='test_loss_small_all_metrics or test_nms_small_dense or test_grouped_schedule_equals_sequential and small or test_golden'
timeout 1500 $CS --tool racecheck --racecheck-report all --error-exitcode 9 --print-limit 20 \
  python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "$RSEL" > gpurun_out/sanitizer_racecheck.txt 2>&1
echo "racecheck rc=$?"; tail -5 gpurun_out/sanitizer_racecheck.txt
timeout 900 $CS --tool synccheck --error-exitcode 9 --print-limit 20 \
  python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "$RSEL" > gpurun_out/sanitizer_synccheck.txt 2>&1
echo "synccheck rc=$?"; tail -5 gpurun_out/sanitizer_synccheck.txt
# the large-segment NMS kernel (cross-CTA protocol in global memory: racecheck only sees shared
# memory, so memcheck + a 3x repeated exactness test is the check here)
timeout 900 $CS --tool memcheck --error-exitcode 9 --print-limit 20 \
  python -m pytest tests/test_gpu_nms.py -m gpu -q -x -p no:cacheprovider -k "by_scores_exact and 20000" > gpurun_out/sanitizer_memcheck_big.txt 2>&1
echo "memcheck big rc=$?"; tail -5 gpurun_out/sanitizer_memcheck_big.txt
