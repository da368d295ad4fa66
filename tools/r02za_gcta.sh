# GCTA f-3 leftovers: the new tests, the existing GCTA tests (unchanged behaviour), the no-alloc test after them, the option's cost.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gcta_diag_gpu.py tests/test_gpu_parity.py tests/test_host_engine.py tests/test_zz_no_alloc.py -m gpu -q -k "gcta or no_alloc or guard or fetch_block" > gpurun_out/r02za_gcta_tests.log 2>&1; echo rc=$?
tail -40 gpurun_out/r02za_gcta_tests.log
timeout 300 python tools/probes/gcta_diag_probe.py 20000 30000 > gpurun_out/r02za_gcta_cost.jsonl 2> gpurun_out/r02za_gcta_cost.err; echo rc=$?
cat gpurun_out/r02za_gcta_cost.jsonl; tail -5 gpurun_out/r02za_gcta_cost.err
