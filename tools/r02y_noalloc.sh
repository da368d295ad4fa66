# The no-alloc entry point's test alone in a fresh process, then the closing run of the round (whole suite -x ends with the same test).
set -x
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02y_noalloc_alone.log 2>&1; echo rc=$?; tail -15 gpurun_out/r02y_noalloc_alone.log
bash tools/r02z_final.sh
