# Round-1 run F: tests after the engine changes (first-slab store, fused LOCO fold, balanced slabs, banded async finish,
# tile / row-slice writers), bench, C4 probe, writer-thread sweep through the CLI.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01f_pytest.log 2>&1; tail -15 gpurun_out/r01f_pytest.log
python bench.py > gpurun_out/r01f_bench_bf16.json 2> gpurun_out/r01f_bench_bf16.err; echo rc=$?; cat gpurun_out/r01f_bench_bf16.json
python tools/config_probe.py C4 C3 --precision bf16 > gpurun_out/r01f_configs.jsonl 2> gpurun_out/r01f_configs.err; echo rc=$?; cat gpurun_out/r01f_configs.jsonl
for wt in 4 8 16; do python tools/cli_configs.py C2 --precision bf16 --writer-threads $wt >> gpurun_out/r01f_cli.jsonl 2>> gpurun_out/r01f_cli.err; done
cat gpurun_out/r01f_cli.jsonl
