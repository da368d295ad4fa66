# Closing 1-GPU run, last of round 2: whole GPU suite, smoke, both bench arms (the SYRK kernel and the C2 launch list are unchanged
# since tools/r02z_final.sh's ncu captures, so those are not repeated).
set -x
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02z_noalloc_alone.log 2>&1; echo rc=$?; tail -2 gpurun_out/r02z_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02z_pytest.log 2>&1; echo rc=$?; tail -4 gpurun_out/r02z_pytest.log
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r02z_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r02z_smoke.log
timeout 400 python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r02z_bench_n1.json 2> gpurun_out/r02z_bench_n1.err; echo rc=$?; cut -c1-400 gpurun_out/r02z_bench_n1.json
timeout 300 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r02z_reference_arm.json 2> gpurun_out/r02z_reference_arm.err; echo rc=$?; cut -c1-300 gpurun_out/r02z_reference_arm.json
