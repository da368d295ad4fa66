# Round-1 run L: parked tail slabs (band-major finish).  Suite, bench, C4/C3 probe, C2 through the CLI.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01l_pytest.log 2>&1; tail -8 gpurun_out/r01l_pytest.log
python bench.py > gpurun_out/r01l_bench.json 2> gpurun_out/r01l_bench.err; echo rc=$?; cat gpurun_out/r01l_bench.json
GGRM_TAIL_SLABS=0 python bench.py --no-cpu > gpurun_out/r01l_bench_tail0.json 2> gpurun_out/r01l_bench_tail0.err; echo rc=$?; cat gpurun_out/r01l_bench_tail0.json
python tools/config_probe.py C4 C3 C1 --precision bf16 > gpurun_out/r01l_configs.jsonl 2> gpurun_out/r01l_configs.err; echo rc=$?; cat gpurun_out/r01l_configs.jsonl
python tools/cli_configs.py C2 --precision bf16 > gpurun_out/r01l_cli.jsonl 2> gpurun_out/r01l_cli.err; cat gpurun_out/r01l_cli.jsonl
