# Round-1 (second half) measurement run: tests, bench (fp16 + bf16), launch list and full SYRK capture.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01b_pytest.log 2>&1; tail -3 gpurun_out/r01b_pytest.log
python bench.py > gpurun_out/r01b_bench_fp16.json 2> gpurun_out/r01b_bench_fp16.err; echo rc=$?
python bench.py --precision bf16 --no-cpu > gpurun_out/r01b_bench_bf16.json 2> gpurun_out/r01b_bench_bf16.err; echo rc=$?
export GGRM_BENCH_M=131072
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01b_small.json 2>&1; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01b_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01b_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:syrk_tc -s 2 -c 1 -o gpurun_out/r01b_syrk python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01b_ncu2.log 2>&1
ls -la gpurun_out/
