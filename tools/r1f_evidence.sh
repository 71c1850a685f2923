# round-1 final evidence run (one B200): tests, bench both arms, ncu launch list + full capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r1f_pytest.log 2>&1; tail -3 gpurun_out/r1f_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1f_smoke.log 2>&1; tail -1 gpurun_out/r1f_smoke.log
python bench.py > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err; echo bench rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1f_bench_ref.json 2>> gpurun_out/r1f_bench.err; echo ref rc=$?
Q="--steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-extras"
python bench.py $Q > gpurun_out/r1f_quick.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1f_launches.csv python bench.py $Q > gpurun_out/r1f_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"loss_dense|loss_rows|loss_finish|nms_filter|decode_kernel|nms_segment|assign" -c 9 -o gpurun_out/r1f_full python bench.py $Q > gpurun_out/r1f_ncu2.log 2>&1
ls -la gpurun_out | tail -12
