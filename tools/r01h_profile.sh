# Round-1 run H: tests (loaders, assoc chunk), then the profile set for profiles/: plain run, launch list, full capture of
# one steady-state SYRK launch (not the banded final slab), at the bench's own settings on a shortened SNP axis.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01h_pytest.log 2>&1; tail -5 gpurun_out/r01h_pytest.log
export GGRM_BENCH_M=131072
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01h_small.json 2> gpurun_out/r01h_small.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01h_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01h_ncu1.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:syrk_tc -s 3 -c 1 -o gpurun_out/r01h_syrk python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01h_ncu2.log 2>&1; echo rc=$?
ls -la gpurun_out/ | grep r01h
