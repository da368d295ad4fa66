# Round-1 run E: GPU tests (new tile / row-slice writers), named configs through the CLI with real files on tmpfs.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r01e_pytest.log 2>&1; tail -15 gpurun_out/r01e_pytest.log
python tools/cli_configs.py C1 C3 C2 C4 --precision bf16 > gpurun_out/r01e_cli.jsonl 2> gpurun_out/r01e_cli.err; echo rc=$?
cat gpurun_out/r01e_cli.jsonl; tail -5 gpurun_out/r01e_cli.err
