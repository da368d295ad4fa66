# expand16 with byte-permute lookups + repack word fast path: full GPU suite, smoke, bench, and a full capture of the prep kernels.
set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/s26d_pytest.log 2>&1; echo rc=$?; tail -4 gpurun_out/s26d_pytest.log
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/s26d_smoke.log 2>&1; echo rc=$?; cat gpurun_out/s26d_smoke.log
timeout 200 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu > gpurun_out/s26d_bench.json 2> gpurun_out/s26d_bench.err; echo rc=$?; cat gpurun_out/s26d_bench.json
export GGRM_BENCH_M=131072
timeout 100 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/s26d_small.json 2> gpurun_out/s26d_small.err; echo rc=$?
timeout 120 ncu --set full --clock-control none --import-source on -k regex:'repack_kernel|count_stats_kernel|expand16_kernel' -s 6 -c 3 -o gpurun_out/s26d_prep python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/s26d_ncu_prep.log 2>&1; echo rc=$?
