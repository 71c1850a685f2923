#!/bin/bash
# r2g ncu evidence (one GPU, ncu only — no sanitizer in the same call): launch list of a short
# bench run, then one --set full capture of the kernels of the step.  Each ncu run directly follows
# a plain run of the same command line that exited 0.
mkdir -p gpurun_out
Q="--steps 3 --warmup 3 --windows 1 --no-e2e --no-cpu-baseline --no-extras --no-strong"
python bench.py $Q > gpurun_out/r2g_quick.json 2> gpurun_out/r2g_quick.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2g_launches.csv python bench.py $Q > gpurun_out/r2g_ncu1.log 2>&1
echo "launch list rc=$?"
python bench.py $Q > gpurun_out/r2g_quick2.json 2>> gpurun_out/r2g_quick.err && \
ncu --set full --clock-control none --import-source on -k regex:"loss_dense|loss_rows|loss_finish|nms_filter|decode_kernel|nms_segment|assign_slots|assign_emit" -c 16 -o gpurun_out/r2g_full python bench.py $Q > gpurun_out/r2g_ncu2.log 2>&1
echo "full capture rc=$?"
# the large-segment NMS kernel on one group of 65 536 rows
python tools/nms_big_once.py > gpurun_out/r2g_big_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"nms_big" -c 1 -o gpurun_out/r2g_big python tools/nms_big_once.py > gpurun_out/r2g_ncu3.log 2>&1
echo "nms_big capture rc=$?"
ls -la gpurun_out | tail -12
