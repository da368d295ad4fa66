#!/bin/bash
# r01x: full GPU suite + smoke + default bench after the genotype-export row
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01x_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r01x_pytest.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r01x_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r01x_smoke.log
timeout 600 python bench.py > gpurun_out/r01x_bench_n1.json 2> gpurun_out/r01x_bench_n1.err; echo "bench rc=$?"; cat gpurun_out/r01x_bench_n1.json
