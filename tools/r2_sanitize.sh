#!/bin/bash
# tools/r2_sanitize.sh <memcheck|racecheck|synccheck>: ONE compute-sanitizer tool per gpurun call
# (B200_PROFILING.md) over the small-input GPU tests.  memcheck: the whole
# "small or golden" selection incl. the plain-C ABI program and a large-segment NMS case;
# racecheck / synccheck: the kernels with hand-rolled shared-memory protocols (loss_finish merged
# stores, nms_segment / nms_big aliased shared memory, assign_emit's EMPTY-slot invariant).
# Summary -> gpurun_out/sanitizer_<tool>.txt (copied to profiles/ by hand).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
if [ "$TOOL" = memcheck ]; then
  SEL='small or golden or abi_smoke or (by_scores_exact and 20000)'
  EXTRA=""
else
  SEL='test_loss_small_all_metrics or test_nms_small_dense or (test_grouped_schedule_equals_sequential and small) or golden or (by_scores_exact and 5000)'
  EXTRA=""
  [ "$TOOL" = racecheck ] && EXTRA="--racecheck-report all"
fi
timeout 2400 $CS --tool $TOOL $EXTRA --error-exitcode 9 --print-limit 30 \
  python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "$SEL" > gpurun_out/sanitizer_$TOOL.txt 2>&1
echo "$TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" gpurun_out/sanitizer_$TOOL.txt | tail -8
