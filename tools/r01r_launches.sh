set -x
export GGRM_BENCH_M=131072
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01r_small.json 2> gpurun_out/r01r_small.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01r_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01r_ncu1.log 2>&1; echo rc=$?
