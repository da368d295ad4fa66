set -x
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_host_engine.py tests/test_gcta_diag_gpu.py tests/test_reference_differential.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02zj_tests.log 2>&1; echo rc=$?; tail -6 gpurun_out/r02zj_tests.log
