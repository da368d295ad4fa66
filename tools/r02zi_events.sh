set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_host_engine.py -m gpu -q -x -k "observer or cli_is_a_drop_in or engine_files or loco_one_pass" > gpurun_out/r02zi_tests.log 2>&1; echo rc=$?; tail -30 gpurun_out/r02zi_tests.log
