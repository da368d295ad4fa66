# C++ engine's band-pipelined SNP-shard path on one GPU (ranks share the device: GELEX_B200_OVERSUBSCRIBE in the tests).
set -x
timeout 300 python -m pytest tests/test_host_engine.py tests/test_multiproc_gpu.py tests/test_sharding.py -m gpu -x -q > gpurun_out/s26_pytest.log 2>&1; echo rc=$?; tail -30 gpurun_out/s26_pytest.log
