# Closing run of the session (one GPU): full GPU suite, smoke, both bench arms as the driver calls them, then full ncu captures of
# the HBM-bound kernels beside the SYRK (repack / count_stats / expand16 on the prep stream, pack_rows on the read-out stream).
# Device code of the SYRK is unchanged since r01z (its launch list and capture stay valid).
set -x
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/s26c_pytest.log 2>&1; echo rc=$?; tail -4 gpurun_out/s26c_pytest.log
timeout 120 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/s26c_smoke.log 2>&1; echo rc=$?; cat gpurun_out/s26c_smoke.log
timeout 200 python bench.py --impl reference --gpus 1 --steps 3 --warmup 3 > gpurun_out/s26c_ref.json 2> gpurun_out/s26c_ref.err; echo rc=$?; tail -c 300 gpurun_out/s26c_ref.json
timeout 300 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/s26c_bench.json 2> gpurun_out/s26c_bench.err; echo rc=$?; cat gpurun_out/s26c_bench.json
export GGRM_BENCH_M=131072
timeout 120 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/s26c_small.json 2> gpurun_out/s26c_small.err; echo rc=$?
timeout 200 ncu --set full --clock-control none --import-source on -k regex:'repack_kernel|count_stats_kernel|expand16_kernel' -s 6 -c 6 -o gpurun_out/s26c_prep python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/s26c_ncu_prep.log 2>&1; echo rc=$?
timeout 200 ncu --set full --clock-control none --import-source on -k regex:pack_rows_kernel -s 8 -c 3 -o gpurun_out/s26c_pack python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/s26c_ncu_pack.log 2>&1; echo rc=$?
ls -la gpurun_out/ | grep s26c
