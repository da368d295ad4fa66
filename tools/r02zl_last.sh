# the in-tree libraries as rebuilt by __graft_entry__.build(): smoke, the no-alloc test alone, one short bench line
set -x
mkdir -p gpurun_out
timeout 60 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r02zl_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r02zl_smoke.log
timeout 120 python -m pytest tests/test_zz_no_alloc.py tests/test_host_engine.py -m gpu -q -k "no_alloc or observer or cli_is_a_drop_in" > gpurun_out/r02zl_tests.log 2>&1; echo rc=$?; tail -2 gpurun_out/r02zl_tests.log
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu --no-wall > gpurun_out/r02zl_bench.json 2> gpurun_out/r02zl_bench.err; echo rc=$?; cut -c1-330 gpurun_out/r02zl_bench.json
