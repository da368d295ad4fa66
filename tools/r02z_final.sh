# Closing 1-GPU run of round 2: whole GPU suite, smoke, both bench arms, launch list and one full capture of a steady-state SYRK launch
# (shortened SNP axis, operand split forced off so that the launches are the ones the 500 000-SNP bench makes).
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02z_pytest.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02z_pytest.log
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r02z_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r02z_smoke.log
timeout 400 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r02z_bench_n1.json 2> gpurun_out/r02z_bench_n1.err; echo rc=$?; cat gpurun_out/r02z_bench_n1.json
timeout 300 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r02z_reference_arm.json 2> gpurun_out/r02z_reference_arm.err; echo rc=$?; cat gpurun_out/r02z_reference_arm.json
export GGRM_BENCH_M=131072 GGRM_OPERAND_SPLIT=0
timeout 200 python bench.py --steps 1 --warmup 1 --no-cpu --no-wall > gpurun_out/r02z_small.json 2> gpurun_out/r02z_small.err; echo rc=$?
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02z_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-wall > gpurun_out/r02z_ncu1.log 2>&1; echo rc=$?
timeout 300 ncu --set full --clock-control none --import-source on -k regex:syrk_tc -s 4 -c 1 -o gpurun_out/r02z_syrk python bench.py --steps 1 --warmup 1 --no-cpu --no-wall > gpurun_out/r02z_ncu2.log 2>&1; echo rc=$?
ls -la gpurun_out | grep r02z
