# Round-1 closing run on one GPU: suite, smoke, both bench arms as the driver calls them, then the profile set of the final
# code (launch list of a shortened bench run + one full capture of a steady-state SYRK launch).
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01z_pytest.log 2>&1; tail -4 gpurun_out/r01z_pytest.log
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r01z_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r01z_smoke.log
python bench.py --impl reference --gpus 1 --steps 3 --warmup 3 > gpurun_out/r01z_ref.json 2> gpurun_out/r01z_ref.err; echo rc=$?; tail -c 600 gpurun_out/r01z_ref.json
python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r01z_bench.json 2> gpurun_out/r01z_bench.err; echo rc=$?; cat gpurun_out/r01z_bench.json
export GGRM_BENCH_M=131072
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01z_small.json 2> gpurun_out/r01z_small.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01z_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01z_ncu1.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:syrk_tc -s 3 -c 1 -o gpurun_out/r01z_syrk python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r01z_ncu2.log 2>&1; echo rc=$?
ls -la gpurun_out/ | grep r01z
