# Round-1 closing run on one GPU: suite, smoke, both bench arms as the driver calls them.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01q_pytest.log 2>&1; tail -6 gpurun_out/r01q_pytest.log
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r01q_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r01q_smoke.log
python bench.py --impl reference --gpus 1 --steps 3 --warmup 3 > gpurun_out/r01q_ref.json 2> gpurun_out/r01q_ref.err; echo rc=$?; tail -c 700 gpurun_out/r01q_ref.json
python bench.py --gpus 1 --steps 3 --warmup 3 > gpurun_out/r01q_bench.json 2> gpurun_out/r01q_bench.err; echo rc=$?; cat gpurun_out/r01q_bench.json
